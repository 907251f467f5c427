#!/usr/bin/env python
"""Time the fused kernel of several builds of the library (launch-bound / tuning variants).
    python tools/variant_bench.py build/variants/*.so
Each library is loaded in its own process; prints ms per launch and env-steps/s.
Environment: VB_ENVS, VB_INNER, VB_DIFF, VB_MAXSTEPS, VB_RING (rollout ring slots), VB_MERGED (0/1)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys
sys.path.insert(0, %r)
from gym_eldorado_b200 import capi, layout
L = capi.load(sys.argv[1])
n, inner = int(sys.argv[2]), int(sys.argv[3])
diff, max_steps = int(sys.argv[4]), int(sys.argv[5])
v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 12345, 0, 0, None, L)
ring, merged = int(sys.argv[6]), bool(int(sys.argv[7]))
v.reset(12345, 4, 3, diff, max_steps)
if ring: v.rollout_attach(ring)
for _ in range(3): v.run_fused(s, inner, merged)
ms = []
for _ in range(8):
    v.run_fused(s, inner, merged); ms.append(v.last_kernel_ms())
ms.sort(); m = ms[len(ms)//2]
print("%%-40s %%8.3f ms/launch  %%.3e env-steps/s" %% (sys.argv[1].split("/")[-1], m, n*inner/(m*1e-3)))
''' % ROOT

if __name__ == "__main__":
    libs = [a for a in sys.argv[1:] if a.endswith(".so")]
    n = os.environ.get("VB_ENVS", "65536"); inner = os.environ.get("VB_INNER", "256")
    diff = os.environ.get("VB_DIFF", "1"); ms = os.environ.get("VB_MAXSTEPS", "100000")
    ring = os.environ.get("VB_RING", "0"); merged = os.environ.get("VB_MERGED", "0")
    for lib in libs:
        subprocess.run([sys.executable, "-c", CHILD, os.path.abspath(lib), n, inner, diff, ms, ring, merged])
