#!/usr/bin/env python
"""Where the time of the packed host loop goes: each call alone, and the step call with some of its result
arrays left out.   python tools/packed_io_bench.py [lib.so]    env: VB_ENVS, VB_ITERS"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gym_eldorado_b200 import capi, layout
L = capi.load(sys.argv[1]) if len(sys.argv) > 1 else capi.lib()
n = int(os.environ.get("VB_ENVS", "65536")); it = int(os.environ.get("VB_ITERS", "200"))


def pinned(nbytes, dt):
    return torch.zeros(nbytes, dtype=torch.uint8).pin_memory().numpy().view(dt)


v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 12345, 0, 0, None, L)
v.reset(12345, 4, 3, 1, 100000)
pm = pinned(n * 16, layout.packed_mask_dtype); pa = pinned(n * 8, layout.packed_action_dtype)
rew = pinned(n * 16, np.float32).reshape(n, 4); done = pinned(n, np.bool_); agent = pinned(n, np.uint8)
v.packed_masks(out=pm)


def timed(name, fn):
    for _ in range(10): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(it): fn()
    torch.cuda.synchronize(); us = (time.perf_counter() - t0) * 1e6 / it
    print("%-58s %8.1f us" % (name, us)); return us


a = timed("sample_packed_host_io", lambda: s.sample_packed_host_io(pm, pa))
b = timed("step_packed_host_io (masks, rewards, dones, agents)", lambda: v.step_packed_host_io(pa, pm, rew, done, agent))
timed("step_packed_host_io (masks, dones, agents)", lambda: v.step_packed_host_io(pa, pm, None, done, agent))
timed("step_packed_host_io (masks)", lambda: v.step_packed_host_io(pa, pm, None, None, None))
timed("step_packed_host_io (nothing back)", lambda: v.step_packed_host_io(pa, None, None, None, None))
def loop():
    s.sample_packed_host_io(pm, pa); v.step_packed_host_io(pa, pm, rew, done, agent)
c = timed("the loop: sample + step", loop)
print("loop %.3e env-steps/s" % (n / c * 1e6))
