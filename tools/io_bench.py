#!/usr/bin/env python
"""Time the two host-buffer entry points separately (pinned arrays: zero-copy kernels; with
COG_IO_ZEROCOPY=0: staged copy pipeline).   python tools/io_bench.py [lib.so]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gym_eldorado_b200 import capi, layout
L = capi.load(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].endswith(".so") else capi.lib()
n = int(os.environ.get("VB_ENVS", "65536")); it = int(os.environ.get("VB_ITERS", "50"))
v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 12345, 0, 0, None, L)
v.reset(12345, 4, 3, 1, 100000)
pin = lambda nbytes: torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
tm, ta, tr, td, tg = pin(n * 128), pin(n * 64), pin(n * 16), pin(n), pin(n)
masks = tm.numpy().view(layout.action_mask_dtype); actions = ta.numpy().view(layout.action_dtype)
rew = tr.numpy().view(np.float32).reshape(n, 4); done = td.numpy().view(np.bool_); agent = tg.numpy()
v.read(layout.BUF_SELECTED_MASKS, masks)
for mode in ("1", "0"):
    os.environ["COG_IO_ZEROCOPY"] = mode
    for _ in range(5):
        s.sample_host_io(masks, actions); v.step_host_io(actions, masks, rew, done, agent)
    t_s = t_v = 0.0
    for _ in range(it):
        t0 = time.perf_counter(); s.sample_host_io(masks, actions); t1 = time.perf_counter()
        v.step_host_io(actions, masks, rew, done, agent); t2 = time.perf_counter()
        t_s += t1 - t0; t_v += t2 - t1
    print("zero_copy=%s  sample_host_io %7.1f us  step_host_io %7.1f us  total %7.1f us  %.3e env-steps/s" % (
        mode, t_s / it * 1e6, t_v / it * 1e6, (t_s + t_v) / it * 1e6, n * it / (t_s + t_v)))
    # variants of the step call: without the mask output, without any output
    for label, args in (("no masks out", (None, rew, done, agent)), ("no outputs", (None, None, None, None))):
        t0 = time.perf_counter()
        for _ in range(it):
            v.step_host_io(actions, *args)
        print("    step_host_io, %-13s %7.1f us" % (label, (time.perf_counter() - t0) / it * 1e6))
