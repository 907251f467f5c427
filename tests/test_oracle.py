"""CPU: pin the oracle (oracle/cog_oracle.c) — against the golden vectors minted from the
unmodified reference, against the reference itself where oracle/_ref was built, and
against the behaviours the reference's own doctests pin
(src/tests/test_environment.cpp)."""
import numpy as np
import pytest

import golden_util
import parity
from gym_eldorado_b200 import layout
from gym_eldorado_b200.layout import payload_bytes as pb
from oracle import port, refshim


@pytest.mark.parametrize("name", sorted(golden_util.TRAJECTORIES))
def test_oracle_matches_golden_trajectory(oracle_cls, name):
    cfg = golden_util.TRAJECTORIES[name]
    o = oracle_cls(cfg[0], cfg[1])
    golden_util.check_trajectory(name, o, lambda: parity.oracle_snapshot(o), parity.merged_masks)
    o.close()


def test_oracle_matches_golden_boards(oracle_cls):
    golden_util.check_boards(lambda n: oracle_cls(n, 0), parity.oracle_snapshot)


def test_prng_known_answers():
    # minstd_rand0: the 10000th output from seed 1 is 1043618065 (C++ standard, [rand.predef])
    st = port.minstd_stream(1, 10000)
    assert st[-1] == 1043618065
    assert port.minstd_stream(0, 3) == port.minstd_stream(1, 3)  # seed 0 -> state 1 (SURVEY A#14)
    assert port.minstd_stream(2147483647, 3) == port.minstd_stream(1, 3)


@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("n,seed,difficulty,max_steps,steps,regime", [
    (16, 777, layout.EASY, 100000, 1200, "A"),
    (64, 777, layout.EASY, 100000, 1200, "M"),
    (64, 777, layout.HARD, 60, 1200, "M"),
    (64, 12345, layout.MEDIUM, 30, 1200, "M"),
    (16, 31, layout.HARD, 5, 600, "M"),
])
def test_oracle_matches_compiled_reference(oracle_cls, n, seed, difficulty, max_steps, steps, regime):
    r = refshim.RefVec(n, seed)
    o = oracle_cls(n, seed)
    r.reset(seed, 4, 3, difficulty, max_steps)
    o.reset(seed, 4, 3, difficulty, max_steps)
    for s in range(steps):
        if regime == "A":
            r.sample(); o.sample()
        else:
            m = r.merged_masks()
            r.sample(m); o.sample(m)
        r.step(); o.step()
        bad = parity.first_mismatch(parity.oracle_snapshot(o), parity.oracle_snapshot(r))
        assert bad is None, "step %d: %s" % (s, bad)
    r.close(); o.close()


@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("difficulty", [0, 1, 2])
def test_oracle_boards_match_compiled_reference(difficulty):
    for seed in range(100, 400):
        for n_pieces in (1, 3, 5):
            if difficulty == 0 and n_pieces > 3:
                continue
            a = refshim.map_generate(seed, n_pieces, difficulty)
            b = port.map_generate(seed, n_pieces, difficulty)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2]), (seed, n_pieces)


# ---- what the reference's doctests pin (src/tests/test_environment.cpp) ----
def test_truncation_length(oracle_cls):
    """:83-103 — a random rollout with max_steps=100 ends with info.total_length == 100."""
    o = oracle_cls(4, 7)
    o.reset(7, 4, 3, layout.EASY, 100)
    seen = 0
    for _ in range(3000):
        o.sample(); o.step()
        for i in np.flatnonzero(o.dones):
            assert o.infos["total_length"][i] == 100
            seen += 1
    assert seen >= 4
    o.close()


def test_win_is_zero_sum(oracle_cls):
    """:106-131 — n_pieces=1 games end by a win; returns sum to zero."""
    o = oracle_cls(16, 42)
    o.reset(54321, 4, 1, layout.EASY, 100000)
    wins = 0
    for _ in range(6000):
        o.sample(parity.merged_masks(parity.oracle_snapshot(o)))
        o.step()
        for i in np.flatnonzero(o.dones):
            r = o.rewards[i]
            assert r.sum() == 0 and r.max() > 0
            assert o.infos["total_length"][i] < 100000
            wins += 1
    assert wins > 0
    o.close()


def test_map_generation_failure(oracle_cls):
    """:68-80 — reset(124,3,3,EASY) works, reset(123,3,4,EASY) throws."""
    o = oracle_cls(1, 0)
    o.reset(124, 3, 3, layout.EASY, 100)
    with pytest.raises(port.MapGenerationError):
        o.reset(123, 3, 4, layout.EASY, 100)
    o.close()


def test_seed_determinism(oracle_cls):
    """:14-65 — same seed, same board; reset() keeps the parameters."""
    a, b = oracle_cls(8, 0), oracle_cls(8, 0)
    a.reset(123, 3, 8, layout.MEDIUM, 200)
    b.reset(123, 3, 8, layout.MEDIUM, 200)
    assert np.array_equal(pb(a.observations), pb(b.observations))
    a.close(); b.close()


# ---- round 2: the remaining links of the chain on the CPU side -------------------------------
@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("difficulty", [1, 2])
def test_oracle_boards_many_pieces_match_compiled_reference(difficulty):
    """boards of up to 8 travel pieces (they reach the edge of the 48x48 grid), on an
    oversized observation buffer so that the reference's writes stay in bounds"""
    for seed in range(5000, 5120):
        for n_pieces in (4, 6, 7, 8):
            a = refshim.map_generate(seed, n_pieces, difficulty, rows=96)
            b = port.map_generate(seed, n_pieces, difficulty, rows=96)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2]), (seed, n_pieces)
            assert np.array_equal(a[1][:6], b[1][:6]), (seed, n_pieces)


@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("difficulty,max_steps", [(layout.EASY, 100000), (layout.HARD, 40)])
def test_oracle_mid_episode_reset_matches_compiled_reference(oracle_cls, difficulty, max_steps):
    """reset() in the middle of an episode, port against the compiled reference"""
    parity.run_mid_episode_reset(lambda n, seed: oracle_cls(n, seed), refshim.RefVec, 64, 4321, difficulty, max_steps)


@pytest.mark.skipif(not refshim.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_two_players_matches_compiled_reference(oracle_cls):
    """2-player games where the reference is defined (every board in the window drew start piece A,
    src/map.cpp:343-354, SURVEY A#13): the port's "player i starts on hex i" rule == the reference"""
    seeds = parity.two_player_clean_seeds(99000, 8192, layout.HARD, 20, 120, want=12)
    assert len(seeds) >= 6
    for seed in seeds:
        r = refshim.RefVec(1, seed)
        o = oracle_cls(1, seed)
        r.reset(seed, 2, 3, layout.HARD, 20)
        o.reset(seed, 2, 3, layout.HARD, 20)
        for s in range(120):
            m = r.merged_masks()
            r.sample(m); o.sample(m)
            r.step(); o.step()
            bad = parity.first_mismatch(parity.oracle_snapshot(o), parity.oracle_snapshot(r))
            assert bad is None, "seed %d step %d: %s" % (seed, s, bad)
        assert o.hidden(0)[14] == 0
        r.close(); o.close()
