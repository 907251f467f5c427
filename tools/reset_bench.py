#!/usr/bin/env python
"""Time vec.reset() (board generation + decks + map encode) for the three difficulties.
    python tools/reset_bench.py [lib.so]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gym_eldorado_b200 import capi
L = capi.load(sys.argv[1]) if len(sys.argv) > 1 else capi.lib()
n = int(os.environ.get("VB_ENVS", "65536"))
for diff, name in ((0, "EASY"), (1, "MEDIUM"), (2, "HARD")):
    v = capi.VecHandle(n, 0, 0, None, L)
    v.reset(12345, 4, 3, diff, 100000)
    for _ in range(2): v.reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): v.reset()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%-28s %-6s %7.3f ms per batch reset  %.3e resets/s" % (os.path.basename(sys.argv[1]) if len(sys.argv) > 1 else "lib", name, ms, n / ms * 1e3))
    v.close()
