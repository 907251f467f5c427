"""CPU: a bounded, fixed-seed slice of the randomised parity campaign (tools/fuzz_emulated.py) — random
env counts, offsets, players, pieces, difficulties, episode lengths and driving modes; the full campaign is
time-boxed and run by hand (profiles/r02_fuzz_campaign.txt)."""
import os
import random
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import fuzz_emulated as fz  # noqa: E402


def test_fuzz_emulated_kernels_vs_restatement(emu, oracle_cls):
    rng = random.Random(20261105)
    kinds = {}
    for _ in range(36):
        case = fz.draw_case(rng)
        case["steps"] = min(case["steps"], 150)
        try:
            kind = fz.run_case(case, emu, oracle_cls)
        except AssertionError as e:
            raise AssertionError("%r: %s" % (case, e))
        kinds[kind] = kinds.get(kind, 0) + 1
    # the slice is varied enough to mean something: single steps and fused launches both occur
    assert len(kinds) >= 5, kinds


def test_fuzz_restatement_vs_compiled_reference():
    from oracle import refshim
    if not refshim.available():
        pytest.skip("oracle/_ref not built")
    rng = random.Random(5112026)
    kinds = {}
    for _ in range(16):
        case = fz.draw_case(rng)
        case["steps"] = min(case["steps"], 150)
        try:
            kind = fz.run_case_port_vs_reference(case)
        except AssertionError as e:
            raise AssertionError("%r: %s" % (case, e))
        kinds[kind] = kinds.get(kind, 0) + 1
    assert len(kinds) >= 3, kinds
