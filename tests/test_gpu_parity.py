"""GPU parity tests: the CUDA kernels, called through the C ABI, against the oracle.
Bit-exact on every payload byte of every exported buffer (integer/byte work; the only
floats are small multiples of 0.5 and integers, compared exactly)."""
import numpy as np
import pytest

import parity
from gym_eldorado_b200 import capi, layout
from oracle import port
from gym_eldorado_b200.layout import payload_bytes as pb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle_cls(gpu_oracle_cls):
    """every test of this module runs against BOTH checkers (ids [port] / [ref]): the C restatement and
    the compiled unmodified reference; cases the reference cannot run (N it is not instantiated for,
    fewer than 4 players) skip on [ref]"""
    return gpu_oracle_cls


@pytest.mark.parametrize("n,seed,difficulty,max_steps,steps,regime,n_players", [
    (16, 0, layout.EASY, 100000, 1000, "A", 4),        # BASELINE config 1 (README example)
    (256, 12345, layout.EASY, 100000, 2000, "A", 4),   # BASELINE config 2 (asv loop), every step
    (256, 777, layout.EASY, 100000, 1500, "M", 4),     # every branch of step, wins, auto-resets
    (256, 777, layout.HARD, 60, 1500, "M", 4),
    (256, 12345, layout.MEDIUM, 30, 1500, "M", 4),
    (128, 99, layout.HARD, 8, 600, "M", 2),            # BASELINE config 4 shape (2 players)
    (128, 99, layout.HARD, 8, 600, "A", 3),
    (64, 4242, layout.HARD, 8, 400, "M", 1),
    (256, 12345, layout.EASY, 100000, 2000, "S", 4),   # configs[1] through the one-kernel sample + step
    (1024, 4242, layout.HARD, 8, 500, "S", 4),
    (64, 31, layout.MEDIUM, 12, 500, "S", 3),
    (256, 5, layout.EASY, 100000, 3000, "P", 4),       # regime P: the agent's own mask record
    (256, 9, layout.HARD, 20, 2000, "P", 4),
    (256, 3, layout.MEDIUM, 0, 100, "M", 4),           # max_steps = 0: every step ends an episode (App. A#8)
    (256, 3, layout.MEDIUM, 0, 100, "S", 4),
    (256, 3, layout.MEDIUM, 1, 300, "M", 4),
])
def test_trajectory_parity(cuda_lib, oracle_cls, n, seed, difficulty, max_steps, steps, regime, n_players):
    parity.run_parity(cuda_lib, oracle_cls, n, seed, difficulty, max_steps, steps, regime,
                      n_players=n_players)


def test_seed_wraps_and_offsets(cuda_lib, oracle_cls):
    # env seed is uint32(seed + i); a shard with an index offset replays its slice of the job
    parity.run_parity(cuda_lib, oracle_cls, 64, 0xFFFFFFF0, layout.MEDIUM, 50, 300, "M")
    parity.run_parity(cuda_lib, oracle_cls, 64, 31337, layout.MEDIUM, 50, 300, "M", offset=1000)


@pytest.mark.parametrize("difficulty", [layout.EASY, layout.MEDIUM, layout.HARD])
def test_boards_on_reset(cuda_lib, oracle_cls, difficulty):
    """Procedural board generation + 48x48x7 encode: 5 x 1024 boards per difficulty."""
    n = 1024
    for seed, n_pieces in ((1, 3), (3001, 3), (500000, 2), (77, 1), (9000, 1)):
        o = oracle_cls(n, 0)
        c = parity.Candidate(cuda_lib, n, 0)
        try:
            o.reset(seed, 4, n_pieces, difficulty, 100)
        except port.MapGenerationError:
            # some one-piece boards cannot host an end piece: the reference throws, so must we
            with pytest.raises(capi.MapGenerationError):
                c.reset(seed, 4, n_pieces, difficulty, 100)
            c.close(); o.close()
            continue
        c.reset(seed, 4, n_pieces, difficulty, 100)
        snap = c.snapshot()
        assert parity.first_mismatch(snap, parity.oracle_snapshot(o)) is None
        # reset() without arguments: RNG streams continue (include/vec_environment.h:33-37)
        o.reset()
        c.reset()
        assert parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o)) is None
        c.close()
        o.close()


def test_map_failure_is_reported(cuda_lib):
    """EASY has only 3 travel pieces: n_pieces=4 throws generate_map_failure in the
    reference (src/tests/test_environment.cpp:68-80) -> MapGenerationError here."""
    from gym_eldorado_b200 import capi
    c = parity.Candidate(cuda_lib, 32, 0)
    with pytest.raises(capi.MapGenerationError):
        c.reset(123, 3, 4, layout.EASY, 100)
    c.reset(124, 3, 3, layout.EASY, 100)  # and the handle stays usable
    c.close()


@pytest.mark.parametrize("n,difficulty,max_steps,k", [(1024, layout.MEDIUM, 100000, 300), (1024, layout.HARD, 8, 300),
                                                      # more envs than resident warps: several envs per warp,
                                                      # episode ends parked for the tail kernel
                                                      (12000, layout.HARD, 8, 120), (12000, layout.EASY, 100000, 120)])
def test_fused_equals_single_steps(cuda_lib, oracle_cls, n, difficulty, max_steps, k):
    """The persistent fused kernel == k x (sample(selected masks); step), bit for bit,
    and both == the oracle."""
    seed = 2024
    o = oracle_cls(n, seed)
    o.reset(seed, 4, 3, difficulty, max_steps)
    for _ in range(k):
        o.sample()
        o.step()
    a = parity.Candidate(cuda_lib, n, seed)
    a.reset(seed, 4, 3, difficulty, max_steps)
    for chunk in (1, 7, k - 8):
        a.vec.run_fused(a.smp, chunk)
    b = parity.Candidate(cuda_lib, n, seed)
    b.reset(seed, 4, 3, difficulty, max_steps)
    for _ in range(k):
        b.sample()
        b.step()
    sa, sb = a.snapshot(), b.snapshot()
    assert parity.first_mismatch(sa, sb) is None
    assert parity.first_mismatch(sa, parity.oracle_snapshot(o)) is None
    st = a.vec.stats()
    assert st["env_steps"] == n * k
    assert st["resets"] == n + st["episodes"]
    a.close(); b.close(); o.close()


def test_full_size_properties(cuda_lib):
    """BASELINE config 3 size (65 536 envs, MEDIUM): properties that need no oracle run.
    Shards replay the same envs; masks stay self-consistent; stats add up."""
    n, seed, k = 65536, 12345, 200
    whole = parity.Candidate(cuda_lib, n, seed)
    whole.reset(seed, 4, 3, layout.MEDIUM, 100000)
    whole.vec.run_fused(whole.smp, k)
    w = whole.snapshot()
    # a shard [40000, 40000+4096) of the same job reproduces its slice exactly
    off, m = 40000, 4096
    part = parity.Candidate(cuda_lib, m, seed, offset=off)
    part.reset(seed, 4, 3, layout.MEDIUM, 100000)
    part.vec.run_fused(part.smp, k)
    p = part.snapshot()
    for name in ("observations", "selected_action_masks", "infos", "actions"):
        assert np.array_equal(pb(w[name][off:off + m]), pb(p[name]))
    for name in ("rewards", "dones", "agent_selection"):
        assert np.array_equal(w[name][off:off + m], p[name])
    obs = w["observations"]
    # card conservation: regime A never buys or removes, so every deck still has 8 cards
    decks = obs["player_data"]["obs"]
    total = sum(decks[f].astype(np.int64).sum(axis=2) for f in ("draw", "hand", "active", "played", "discard"))
    assert (total == 8).all()
    assert (obs["shared"]["map"][..., 0] == 0).all()          # feature 0 is never written
    assert (w["selected_action_masks"]["play"][:, 0]).all()    # index 0 always valid
    assert whole.vec.stats()["env_steps"] == n * k
    whole.close(); part.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,difficulty,max_steps,n_players", [(1024, layout.EASY, 60, 4), (9000, layout.HARD, 40, 4),
                                                             (256, layout.MEDIUM, 100000, 3)])
def test_fused_merged_masks(cuda_lib, oracle_cls, n, difficulty, max_steps, n_players):
    """The fused kernel sampling from the merged legal mask (moves, purchases, removals,
    wins, auto-resets; several envs per warp) == the oracle, bit for bit."""
    st = parity.run_fused_merged_parity(cuda_lib, oracle_cls, n, 31337, difficulty, max_steps, (1, 60, 300),
                                        n_players=n_players)
    assert st["episodes"] > 0 and st["travelled_hexes"] > 0
    assert st["env_steps"] == n * 361


@pytest.mark.gpu
@pytest.mark.parametrize("max_steps", [0, 1, 2])
def test_fused_every_step_ends_an_episode(cuda_lib, oracle_cls, max_steps):
    """max_steps = 0 / 1 / 2: the hot kernel parks (almost) every env at once and the tail
    kernel does the episode ends, board generation and the remaining iterations"""
    st = parity.run_fused_merged_parity(cuda_lib, oracle_cls, 1024, 11, layout.MEDIUM, max_steps, (1, 3, 20))
    assert st["env_steps"] == 1024 * 24 and st["episodes"] == st["truncations"] > 0
    if max_steps == 0:
        assert st["episodes"] == st["env_steps"]


@pytest.mark.gpu
@pytest.mark.parametrize("n,regime,pinned", [(20000, "A", False), (9001, "M", False), (20000, "A", True),
                                             (9001, "M", True)])
def test_host_io_pipeline(cuda_lib, oracle_cls, n, regime, pinned):
    """sample / step through HOST buffers with chunked copy/compute overlap (several chunks,
    a ragged last one) == the oracle at every step."""
    assert parity.run_host_io_parity(cuda_lib, oracle_cls, n, 99, layout.HARD, 10, 150, regime, pinned=pinned) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,merged,difficulty,max_steps,n_players,chunks,n_slots", [
    (6000, False, layout.HARD, 8, 4, (3, 30, 20), 32),      # several envs per warp, ring wraps, tail kernel
    (3000, True, layout.EASY, 60, 4, (40, 60), 128),        # moves, purchases, removals
    (64, True, layout.EASY, 100000, 4, (900, 600), 2048),   # wins: non-zero rewards in the ring
    (500, True, layout.HARD, 12, 2, (25, 48), 16),
])
def test_rollout_ring_and_compact_obs(cuda_lib, oracle_cls, n, merged, difficulty, max_steps, n_players, chunks,
                                      n_slots):
    """the rollout ring written by the fused kernels, cog_vec_compact_obs
    and the one-byte hex grid == records restated from the oracle, byte for byte."""
    dones, wins = parity.run_rollout_parity(cuda_lib, oracle_cls, n, 4711, difficulty, max_steps, chunks, n_slots,
                                            merged, n_players=n_players)
    assert dones > 0
    if max_steps > 100:
        assert wins > 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,regime,pinned", [(16384, "A", True), (4096, "M", False), (1024, "M", True)])
def test_packed_host_io(cuda_lib, oracle_cls, n, regime, pinned):
    """The packed host-I/O entry points (16-byte masks, 8-byte actions: 66 bytes per env-step over PCIe)
    == the oracle at every step, pageable and pinned arrays."""
    assert parity.run_packed_io_parity(cuda_lib, oracle_cls, n, 99, layout.HARD, 10, 120, regime, pinned=pinned) > 0
