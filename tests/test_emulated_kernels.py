"""CPU: the kernels' game logic, compiled for the host under the lock-step warp
emulator (tests/emu — test infrastructure, never loaded by the package), against the
oracle.  Same driver as the GPU parity tests, smaller sizes."""
import pytest

import parity
from gym_eldorado_b200 import capi, layout


@pytest.mark.parametrize("n,seed,difficulty,max_steps,steps,regime,n_players", [
    (16, 0, layout.EASY, 100000, 300, "A", 4),
    (16, 777, layout.EASY, 100000, 700, "M", 4),
    (16, 777, layout.HARD, 60, 700, "M", 4),
    (16, 12345, layout.MEDIUM, 30, 700, "M", 4),
    (8, 99, layout.HARD, 8, 300, "M", 2),
    (8, 99, layout.HARD, 8, 300, "A", 3),
    (16, 5, layout.HARD, 8, 500, "S", 4),       # sample + step in one kernel (cog_vec_step_sampled)
    (8, 6, layout.EASY, 100000, 300, "S", 2),
    (16, 5, layout.EASY, 100000, 400, "P", 4),  # regime P: the agent's own mask record (plays S would refuse)
    (16, 9, layout.HARD, 20, 400, "P", 4),
    (16, 3, layout.MEDIUM, 0, 40, "M", 4),      # max_steps = 0: every step ends an episode
    (16, 3, layout.MEDIUM, 1, 60, "A", 4),
])
def test_emulated_trajectory_parity(emu, oracle_cls, n, seed, difficulty, max_steps, steps, regime, n_players):
    parity.run_parity(emu, oracle_cls, n, seed, difficulty, max_steps, steps, regime, n_players=n_players)


def test_emulated_fused_equals_single(emu, oracle_cls):
    n, seed, k = 12, 2024, 120
    o = oracle_cls(n, seed)
    o.reset(seed, 4, 3, layout.HARD, 8)
    for _ in range(k):
        o.sample()
        o.step()
    a = parity.Candidate(emu, n, seed)
    a.reset(seed, 4, 3, layout.HARD, 8)
    for chunk in (1, 0, 19, k - 20):
        a.vec.run_fused(a.smp, chunk)
    assert parity.first_mismatch(a.snapshot(), parity.oracle_snapshot(o)) is None
    st = a.vec.stats()
    assert st["env_steps"] == n * k and st["resets"] == n + st["episodes"] and st["episodes"] > 0
    a.close(); o.close()


def test_emulated_map_failure(emu):
    c = parity.Candidate(emu, 4, 0)
    with pytest.raises(capi.MapGenerationError):
        c.reset(123, 3, 4, layout.EASY, 100)
    c.reset(124, 3, 3, layout.EASY, 100)
    c.close()


def test_emulated_offset_shard(emu, oracle_cls):
    parity.run_parity(emu, oracle_cls, 8, 31337, layout.MEDIUM, 50, 150, "M", offset=1000)


@pytest.mark.parametrize("max_steps", [0, 1])
def test_emulated_fused_every_step_ends_an_episode(emu, oracle_cls, max_steps):
    st = parity.run_fused_merged_parity(emu, oracle_cls, 16, 11, layout.MEDIUM, max_steps, (1, 3, 7))
    assert st["env_steps"] == 16 * 11 and st["episodes"] == st["truncations"] > 0


def test_emulated_fused_merged_masks(emu, oracle_cls):
    """regime M inside the fused kernel: moves, purchases, removals, wins, auto-resets"""
    st = parity.run_fused_merged_parity(emu, oracle_cls, 12, 777, layout.EASY, 60, (1, 0, 37, 400))
    assert st["episodes"] > 0 and st["travelled_hexes"] > 0 and st["cards_added"] > 0
    assert st["env_steps"] == 12 * 438


@pytest.mark.parametrize("zero_copy,regime", [("1", "A"), ("0", "A"), ("1", "M"), (None, "A")])
def test_emulated_host_io(emu, oracle_cls, monkeypatch, zero_copy, regime):
    # both branches of the *_host_io entry points: kernels addressing the caller's arrays directly
    # (what pinned memory can get on the GPU) and the chunked copy/compute pipeline; None = the
    # default, where the library times both on its first nine calls and keeps the faster
    if zero_copy is None:
        monkeypatch.delenv("COG_IO_ZEROCOPY", raising=False)
    else:
        monkeypatch.setenv("COG_IO_ZEROCOPY", zero_copy)
    assert parity.run_host_io_parity(emu, oracle_cls, 10, 5, layout.HARD, 6, 120, regime) > 0


@pytest.mark.parametrize("merged,difficulty,max_steps,n_players,chunks,n_slots", [
    (False, layout.HARD, 8, 4, [5, 40, 33], 64),      # episode ends inside launches: tail kernel fills rewards
    (True, layout.EASY, 100000, 4, [30, 500, 300], 1024),  # moves, purchases, a win (non-zero rewards)
    (True, layout.HARD, 12, 3, [25, 48], 16),          # ring wraps, also within one launch
])
def test_emulated_rollout_and_compact_obs(emu, oracle_cls, merged, difficulty, max_steps, n_players, chunks, n_slots):
    # compact records in the rollout ring and from cog_vec_compact_obs
    dones, wins = parity.run_rollout_parity(emu, oracle_cls, 12, 4711, difficulty, max_steps, chunks, n_slots,
                                            merged, n_players=n_players)
    assert dones > 0
    if max_steps > 100:
        assert wins > 0


def test_emulated_error_behaviour_of_the_additions(emu, oracle_cls):
    """Argument / state errors of the entry points that are not in the reference: they report
    through cog_last_error like the rest of the C ABI and leave the handle usable."""
    c = parity.Candidate(emu, 5, 1)  # 5 envs: a ragged last block
    try:
        with pytest.raises(capi.CogError) as e:
            c.vec.compact_obs()
        assert e.value.code == capi.COG_ERR_NOT_RESET
        with pytest.raises(capi.CogError):
            c.vec.read_rollout()                  # no ring attached
        c.reset(1, 4, 3, layout.EASY, 100000)
        with pytest.raises(capi.CogError) as e:
            c.vec.rollout_attach(0xffffffff)      # n_envs * n_slots must fit 32 bits
        assert e.value.code == capi.COG_ERR_INVALID
        assert c.vec.rollout_info() == (None, 0, 0)
        c.vec.rollout_attach(1)                   # a ring of one slot keeps the latest step
        c.vec.run_fused(c.smp, 7)
        ring = c.vec.read_rollout()
        assert ring.shape == (5, 1) and (ring["t"] == 6).all() and (ring["filled"] == 1).all()
        c.vec.rollout_attach(3)                   # re-attach: counters restart
        assert c.vec.rollout_info()[1:] == (3, 0)
        c.vec.run_fused(c.smp, 0)                 # zero iterations: nothing recorded
        assert c.vec.rollout_info()[2] == 0
        o = oracle_cls(5, 1)
        o.reset(1, 4, 3, layout.EASY, 100000)
        for _ in range(7):
            o.sample(); o.step()
        assert parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o)) is None
        o.close()
    finally:
        c.close()


# ---- round 2: the GPU parity chain (tests/test_gpu_chain.py) at emulator sizes -----------------
def test_emulated_golden_boards(emu):
    """the kernels' board generation + map encode against the fixtures minted from the reference"""
    import golden_util
    golden_util.check_boards(lambda n: parity.Candidate(emu, n, 0), lambda c: c.snapshot())


@pytest.mark.parametrize("name,steps", [("readme_A", 200), ("hard_reset_heavy_M", 250), ("easy_M", 250)])
def test_emulated_golden_trajectory_prefix(emu, name, steps):
    import golden_util
    cfg = golden_util.TRAJECTORIES[name]
    c = parity.Candidate(emu, cfg[0], cfg[1])
    golden_util.check_trajectory(name, c, c.snapshot, parity.merged_masks, steps=steps)
    c.close()


@pytest.mark.parametrize("difficulty,n_pieces", [(layout.MEDIUM, 5), (layout.HARD, 8), (layout.MEDIUM, 8)])
def test_emulated_boards_many_pieces(emu, oracle_cls, difficulty, n_pieces):
    from oracle import port
    for seed in (11, 4000):
        o = oracle_cls(24, 0)
        c = parity.Candidate(emu, 24, 0)
        try:
            o.reset(seed, 4, n_pieces, difficulty, 100)
        except port.MapGenerationError:
            with pytest.raises(capi.MapGenerationError):
                c.reset(seed, 4, n_pieces, difficulty, 100)
        else:
            c.reset(seed, 4, n_pieces, difficulty, 100)
            assert parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o)) is None
        c.close(); o.close()


def test_emulated_mid_episode_reset(emu, oracle_cls):
    parity.run_mid_episode_reset(lambda n, seed: parity.Candidate(emu, n, seed), oracle_cls, 16, 4321, layout.EASY,
                                 100000)


def test_emulated_two_players_vs_reference(emu):
    """the kernels against the compiled reference itself on 2-player games the reference defines"""
    from oracle import refshim
    if not refshim.available():
        pytest.skip("oracle/_ref not built")
    for seed in parity.two_player_clean_seeds(99000, 8192, layout.HARD, 20, 120, want=3):
        r = refshim.RefVec(1, seed)
        c = parity.Candidate(emu, 1, seed)
        r.reset(seed, 2, 3, layout.HARD, 20)
        c.reset(seed, 2, 3, layout.HARD, 20)
        for s in range(120):
            m = r.merged_masks()
            r.sample(m); c.sample(m)
            r.step(); c.step()
            bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(r))
            assert bad is None, "seed %d step %d: %s" % (seed, s, bad)
        c.close(); r.close()


@pytest.mark.parametrize("regime", ["A", "M"])
def test_emulated_packed_host_io(emu, oracle_cls, regime):
    """16-byte masks / 8-byte actions through cog_*_packed_host_io == the oracle at every step"""
    assert parity.run_packed_io_parity(emu, oracle_cls, 12, 99, layout.HARD, 10, 200, regime) > 0


def test_emulated_wide_boards(oracle_cls):
    """-DCOG_WIDE_BOARDS: board generation by the piece list (bg_generate_wide), the exact path for generations
    that pile up more pieces than the fast path has lanes for or whose intermediate board outgrows the bitmap.
    This build sends every board of more than two pieces through it (COG_BG_MAX_PLACED=2): boards, failed
    generations, auto-resets in long trajectories and the case the randomised campaign found (7 pieces, HARD,
    32 pieces piled up in an auto-reset at step 59, env 30) must all equal the restatement."""
    import os
    import subprocess
    from oracle import port
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")
    subprocess.check_call(["make", "-s", "-C", here, os.path.join(here, "libcog_emu_wide.so")])
    wide = capi.load(os.path.join(here, "libcog_emu_wide.so"))
    for difficulty, n_pieces, seed in ((layout.EASY, 3, 1), (layout.MEDIUM, 5, 11), (layout.HARD, 8, 4000),
                                       (layout.EASY, 4, 123)):
        o = oracle_cls(24, 0)
        c = parity.Candidate(wide, 24, 0)
        try:
            o.reset(seed, 4, n_pieces, difficulty, 100)
        except port.MapGenerationError:
            with pytest.raises(capi.MapGenerationError):
                c.reset(seed, 4, n_pieces, difficulty, 100)
        else:
            c.reset(seed, 4, n_pieces, difficulty, 100)
            assert parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o)) is None
        c.close(); o.close()
    assert parity.run_parity(wide, oracle_cls, 16, 777, layout.HARD, 8, 200, "M") > 0
    parity.run_parity(wide, oracle_cls, 40, 68145605, layout.HARD, 0, 62, "M", n_pieces=7)
