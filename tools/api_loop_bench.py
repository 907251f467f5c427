#!/usr/bin/env python
"""The reference's Python loop on device buffers: sample(selected masks); step(actions) — times it
and (under ncu) exposes the two single-step kernels.   python tools/api_loop_bench.py [lib.so]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gym_eldorado_b200 import capi, layout
L = capi.load(sys.argv[1]) if len(sys.argv) > 1 else capi.lib()
n = int(os.environ.get("VB_ENVS", "65536")); it = int(os.environ.get("VB_ITERS", "200"))
v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 12345, 0, 0, None, L)
v.reset(12345, 4, 3, 1, 100000)
mp, ap = v.buffer(layout.BUF_SELECTED_MASKS)[0], s.actions_ptr()
for _ in range(10): s.sample_device(mp); v.step_device(ap)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(it): s.sample_device(mp); v.step_device(ap)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / it
print("%-24s %8.1f us per sample+step  %.3e env-steps/s" % (os.path.basename(sys.argv[1]) if len(sys.argv) > 1 else "lib", us, n / us * 1e6))
