#!/usr/bin/env python
"""Compare a variant build of the library with the in-tree one on the GPU: same seeds, fused
launches with episode ends and several envs per warp, every exported buffer and the statistics.
    python tools/variant_check.py build/variants/NAME.so"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import parity
from gym_eldorado_b200 import capi, layout
ok = True
for n, diff, max_steps, chunks, merged in ((12000, layout.HARD, 8, (5, 300, 77), False),
                                           (12000, layout.MEDIUM, 100000, (400,), False),
                                           (9000, layout.EASY, 40, (250, 100), True)):
    res = []
    for L in (capi.lib(), capi.load(os.path.abspath(sys.argv[1]))):
        c = parity.Candidate(L, n, 31337)
        c.reset(31337, 4, 3, diff, max_steps)
        for k in chunks:
            c.vec.run_fused(c.smp, k, merged=merged)
        res.append((c.snapshot(), c.vec.stats()))
        c.close()
    bad = parity.first_mismatch(res[0][0], res[1][0])
    same = bad is None and res[0][1] == res[1][1]
    ok &= same
    print("n=%d diff=%d max_steps=%d merged=%s: %s %s" % (n, diff, max_steps, merged, "identical" if same else "DIFFERENT",
                                                       bad or (res[0][1], res[1][1]) if not same else ""))
sys.exit(0 if ok else 1)
