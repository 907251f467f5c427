"""GPU: a bounded, fixed-seed slice of the randomised parity campaign (tools/fuzz_emulated.py, --pair cuda-port):
the product library through the C ABI against the C restatement on random env counts, offsets, players, pieces,
difficulties, episode lengths and driving modes (single steps in every regime, fused launches, host and packed
I/O, the rollout ring, a mid-episode reset).  The time-boxed campaign itself: profiles/r02_z_gpu_fuzz.txt."""
import os
import random
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import fuzz_emulated as fz  # noqa: E402


@pytest.mark.gpu
def test_fuzz_cuda_vs_restatement(cuda_lib, oracle_cls):
    rng = random.Random(911)
    kinds = {}
    for _ in range(60):
        case = fz.draw_case(rng)
        try:
            kind = fz.run_case(case, cuda_lib, oracle_cls)
        except AssertionError as e:
            raise AssertionError("%r: %s" % (case, e))
        kinds[kind] = kinds.get(kind, 0) + 1
    assert len(kinds) >= 6, kinds
