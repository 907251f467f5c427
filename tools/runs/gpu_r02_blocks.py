#!/usr/bin/env python
"""A/B of the envs-per-block of k_reset and k_step_single (COG_RESET_WARPS / COG_SINGLE_WARPS with their resident
blocks) on the GPU, no torch: per build, whole-batch resets per second, the one-kernel per-step loop
(cog_vec_step_sampled) and a digest of the results, which must not depend on the block size.
    python tools/runs/gpu_r02_blocks.py build/variants/v0.so build/variants/v1.so ..."""
import hashlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from gym_eldorado_b200 import capi, layout  # noqa: E402

N = int(os.environ.get("AB_ENVS", "65536"))
ITERS = int(os.environ.get("AB_ITERS", "300"))
ref = None
ok = True
for path in sys.argv[1:]:
    L = capi.load(os.path.abspath(path))
    name = os.path.basename(path)
    out = []
    for diff, dname in ((layout.MEDIUM, "MEDIUM"), (layout.HARD, "HARD")):
        v = capi.VecHandle(N, 0, 0, None, L)
        v.reset(12345, 4, 3, diff, 100000)
        for _ in range(3):
            v.reset()
        v.sync()
        t0 = time.perf_counter()
        for _ in range(20):
            v.reset()
        v.sync()
        dt = (time.perf_counter() - t0) / 20
        out.append("%s %.3f ms = %.3e resets/s" % (dname, dt * 1e3, N / dt))
        v.close()
    for diff, max_steps, dname in ((layout.MEDIUM, 100000, "configs[2]"), (layout.HARD, 8, "reset-heavy")):
        v = capi.VecHandle(N, 0, 0, None, L)
        s = capi.SamplerHandle(N, 12345, 0, 0, None, L)
        v.reset(12345, 4, 3, diff, max_steps)
        for _ in range(20):
            v.step_sampled(s)
        v.sync()
        t0 = time.perf_counter()
        for _ in range(ITERS):
            v.step_sampled(s)
        v.sync()
        dt = (time.perf_counter() - t0) / ITERS
        out.append("step loop %s %.1f us = %.3e env-steps/s" % (dname, dt * 1e6, N / dt))
        h = hashlib.sha1()
        for buf in (layout.BUF_SELECTED_MASKS, layout.BUF_REWARDS, layout.BUF_DONES, layout.BUF_INFOS,
                    layout.BUF_AGENT_SELECTION):
            h.update(np.ascontiguousarray(v.read(buf)).tobytes())
        h.update(np.ascontiguousarray(s.read_actions()).tobytes())
        h.update(np.ascontiguousarray(v.compact_obs(merged=True)).tobytes())
        h.update(repr(sorted(v.stats().items())).encode())
        out.append(h.hexdigest()[:12])
        v.close(); s.close()
    d = (out[3], out[5])
    if ref is None:
        ref = d
    ok &= d == ref
    print("%-8s %s" % (name, " | ".join(out)) + ("" if d == ref else "  DIFFERENT"), flush=True)
print("ALL RESULTS IDENTICAL" if ok else "RESULTS DIFFER")
sys.exit(0 if ok else 1)
