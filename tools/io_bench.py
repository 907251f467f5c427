#!/usr/bin/env python
"""Time the host-buffer entry points (pipelined vs serial) — run with COG_IO_CHUNKS=k to vary the chunking."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gym_eldorado_b200 import capi, layout
L = capi.lib(); n = 65536
vec = capi.VecHandle(n, 0, 0, None, L); smp = capi.SamplerHandle(n, 12345, 0, 0, None, L)
vec.reset(12345, 4, 3, 1, 100000)
pin = lambda nbytes: torch.empty(nbytes, dtype=torch.uint8).pin_memory()
hm, ha, hr, hd, hg = pin(n * 128), pin(n * 64), pin(n * 16), pin(n), pin(n)
m, a = hm.numpy().view(layout.action_mask_dtype), ha.numpy().view(layout.action_dtype)
r, d, g = hr.numpy().view(np.float32).reshape(n, 4), hd.numpy().view(np.bool_), hg.numpy()
vec.read(layout.BUF_SELECTED_MASKS, m)
def t(fn, it=40):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(it): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / it * 1e3
print("chunks", os.environ.get("COG_IO_CHUNKS", "8"),
      "sample_io %.3f ms" % t(lambda: smp.sample_host_io(m, a)),
      "step_io %.3f ms" % t(lambda: vec.step_host_io(a, m, r, d, g)),
      "step_io(masks only) %.3f ms" % t(lambda: vec.step_host_io(a, m)),
      "h2d 8MB %.3f ms" % t(lambda: smp.sample_host(m)),
      "d2h 8MB %.3f ms" % t(lambda: vec.read(layout.BUF_SELECTED_MASKS, m)))
