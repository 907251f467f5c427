#!/usr/bin/env python
"""A/B of the ticketed work units (-DCOG_DYNAMIC_ENVS, build/variants/dyn.so) against the static grid stride
(build/variants/base.so) on the GPU, one process, no torch: for every workload the same launches through both
libraries (and several COG_FUSED_ROUNDS settings of the ticketed one), median kernel time of the launch pair
(k_step_fused + k_step_tail) and a digest of the results — which must not depend on how the work was handed out.
    python tools/runs/gpu_r02_ab.py build/variants/base.so build/variants/dyn.so"""
import hashlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
from gym_eldorado_b200 import capi, layout  # noqa: E402

T0 = time.time()
BUDGET = float(os.environ.get("AB_BUDGET_S", "70"))
base, dyn = capi.load(os.path.abspath(sys.argv[1])), capi.load(os.path.abspath(sys.argv[2]))
N0 = int(os.environ.get("AB_ENVS", "65536"))
# name, difficulty, max_steps, merged, inner, players, ring slots, [(library, COG_FUSED_ROUNDS)...]
# (COG_FUSED_ROUNDS only meant something to the first build measured, which could also split an env's steps
#  into rounds of smaller work units: profiles/r02_ab_ticketed_units.txt; the kept build ignores it)
BOTH = [("base", None), ("dyn", None)]
WORK = [
    ("regime A (headline)", layout.MEDIUM, 100000, False, 1024, 4, 0, BOTH + BOTH),
    ("regime M", layout.MEDIUM, 100000, True, 256, 4, 0, BOTH),
    ("regime M, episode ends (EASY, 40)", layout.EASY, 40, True, 256, 4, 0, BOTH),
    ("reset-heavy configs[3] (HARD, 8, 2 players)", layout.HARD, 8, False, 256, 2, 0, BOTH),
    ("reset-heavy, 4 players (HARD, 8)", layout.HARD, 8, False, 256, 4, 0, BOTH),
    ("regime A + rollout ring (16 slots)", layout.MEDIUM, 100000, False, 256, 4, 16, BOTH),
    ("regime M + rollout ring (16 slots)", layout.MEDIUM, 100000, True, 256, 4, 16, BOTH),
    ("configs[4] shard: 131072 envs, HARD", layout.HARD, 100000, False, 256, 4, -131072, BOTH),
]


def digest(v, s):
    h = hashlib.sha1()
    for buf in (layout.BUF_SELECTED_MASKS, layout.BUF_REWARDS, layout.BUF_DONES, layout.BUF_INFOS,
                layout.BUF_AGENT_SELECTION):
        h.update(np.ascontiguousarray(v.read(buf)).tobytes())
    h.update(np.ascontiguousarray(s.read_actions()).tobytes())
    h.update(np.ascontiguousarray(v.compact_obs(merged=True)).tobytes())
    st = v.stats()
    h.update(repr(sorted(st.items())).encode())
    if v.rollout_info()[1]:
        h.update(np.ascontiguousarray(v.read_rollout()).tobytes())
    return h.hexdigest()[:16], st


ok = True
for name, diff, max_steps, merged, inner, players, ring, combos in WORK:
    N = min(-ring, 2 * N0) if ring < 0 else N0
    ring = max(ring, 0)
    print("== %s: %d envs, %d steps per launch" % (name, N, inner), flush=True)
    ref = None
    for which, rounds in combos:
        if time.time() - T0 > BUDGET:
            print("   (time budget reached, remaining combinations skipped)")
            break
        if rounds is None:
            os.environ.pop("COG_FUSED_ROUNDS", None)
        else:
            os.environ["COG_FUSED_ROUNDS"] = rounds
        L = base if which == "base" else dyn
        v = capi.VecHandle(N, 0, 0, None, L)
        s = capi.SamplerHandle(N, 12345, 0, 0, None, L)
        v.reset(12345, players, 3, diff, max_steps)
        if ring:
            v.rollout_attach(ring)
        ms = []
        for it in range(7):
            v.run_fused(s, inner, merged)
            v.sync()
            if it >= 2:
                ms.append(v.last_kernel_ms())
        ms.sort()
        m = ms[len(ms) // 2]
        d, st = digest(v, s)
        if ref is None:
            ref = d
        same = d == ref
        ok &= same
        print("   %-5s rounds=%-5s %8.3f ms  %.3e env-steps/s  digest %s %s  episodes %d" % (
            which, rounds or "auto", m, N * inner / (m * 1e-3), d, "same" if same else "DIFFERENT", st["episodes"]),
            flush=True)
        v.close(); s.close()
# full snapshots (every exported buffer, the 17 KB observations included) on a smaller batch
import parity  # noqa: E402
os.environ.pop("COG_FUSED_ROUNDS", None)
for n, diff, max_steps, chunks, merged, rounds in ((12000, layout.HARD, 8, (5, 300, 77), False, "3"),
                                                   (20000, layout.EASY, 40, (250, 100), True, None)):
    if time.time() - T0 > BUDGET + 15:
        print("full-snapshot check skipped (time budget)")
        break
    res = []
    for L in (base, dyn):
        if rounds and L is dyn:
            os.environ["COG_FUSED_ROUNDS"] = rounds
        else:
            os.environ.pop("COG_FUSED_ROUNDS", None)
        c = parity.Candidate(L, n, 31337)
        c.reset(31337, 4, 3, diff, max_steps)
        for k in chunks:
            c.vec.run_fused(c.smp, k, merged=merged)
        res.append((c.snapshot(), c.vec.stats()))
        c.close()
    bad = parity.first_mismatch(res[0][0], res[1][0])
    same = bad is None and res[0][1] == res[1][1]
    ok &= same
    print("full snapshot n=%d max_steps=%d merged=%s rounds=%s: %s %s" % (
        n, max_steps, merged, rounds or "auto", "identical" if same else "DIFFERENT", bad or ""), flush=True)
print("total %.1f s; %s" % (time.time() - T0, "ALL RESULTS IDENTICAL" if ok else "RESULTS DIFFER"))
sys.exit(0 if ok else 1)
