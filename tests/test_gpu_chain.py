"""GPU parity chain, closed on the GPU itself (VERDICT r1 "next" #1): the CUDA path, through the
C ABI, against
  - the golden fixtures minted from the UNMODIFIED reference (tests/golden/, tools/make_golden.py),
  - the compiled reference itself (oracle/_ref, ids [ref]) and the C restatement (ids [port]),
at the sizes the contract names: BASELINE.md §4's 256 envs x 10 000 steps x 3 difficulties x
regimes A and M, BASELINE.json configs[2] at 65 536 envs, configs[4]'s per-rank offsets, boards of
4..8 pieces, a mid-episode reset(), and 2-player games against the reference where the reference is
defined (every board in the window drew start piece A, SURVEY A#13).  Bit-exact, payload bytes only."""
import numpy as np
import pytest

import golden_util
import parity
from gym_eldorado_b200 import capi, layout
from gym_eldorado_b200.layout import payload_bytes as pb
from oracle import port, refshim

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle_cls(gpu_oracle_cls):
    return gpu_oracle_cls


# ---- (a) golden fixtures from the reference, replayed on the CUDA path ------------------------
@pytest.mark.parametrize("name", sorted(golden_util.TRAJECTORIES))
def test_cuda_matches_golden_trajectory(cuda_lib, name):
    """oracle: tests/golden/traj_<name>.npz (per-step digests + final payload of the compiled reference)"""
    cfg = golden_util.TRAJECTORIES[name]
    c = parity.Candidate(cuda_lib, cfg[0], cfg[1])
    try:
        golden_util.check_trajectory(name, c, c.snapshot, parity.merged_masks)
    finally:
        c.close()


def test_cuda_matches_golden_boards(cuda_lib):
    """oracle: tests/golden/boards.npz (48 boards per difficulty from the compiled reference)"""
    golden_util.check_boards(lambda n: parity.Candidate(cuda_lib, n, 0), lambda c: c.snapshot())


# ---- (b) BASELINE.md §4: 256 envs x 10 000 steps x 3 difficulties x {A, M} --------------------
SMALL = ("selected_action_masks", "rewards", "dones", "agent_selection", "actions")


def _small_mismatch(c, o, merged_dev=None, merged_want=None):
    got = {"selected_action_masks": c.vec.read(layout.BUF_SELECTED_MASKS), "rewards": c.vec.read(layout.BUF_REWARDS),
           "dones": c.vec.read(layout.BUF_DONES), "agent_selection": c.vec.read(layout.BUF_AGENT_SELECTION),
           "actions": c.smp.read_actions()}
    for k in ("selected_action_masks", "actions"):
        if not np.array_equal(pb(got[k]), pb(getattr(o, k))):
            return k
    for k in ("rewards", "dones", "agent_selection"):
        if not np.array_equal(np.asarray(got[k]), np.asarray(getattr(o, k))):
            return k
    return None


@pytest.mark.parametrize("difficulty,max_steps,regime", [
    (layout.EASY, 100000, "A"), (layout.MEDIUM, 100000, "A"), (layout.HARD, 100000, "A"),
    (layout.EASY, 100000, "M"),   # long episodes that end by a win
    (layout.MEDIUM, 60, "M"), (layout.HARD, 30, "M"),
])
def test_baseline_gate_256x10000(cuda_lib, oracle_cls, difficulty, max_steps, regime):
    """Every step: sampled actions, selected masks, rewards, dones, agent_selection (and, in regime M, the
    device-merged legal mask against the host merge of the oracle's records); every 250th step and the last:
    every payload byte of every exported buffer (ObsData, Info included)."""
    import torch
    n, seed, steps = 256, 12345, 10000
    o = oracle_cls(n, seed)
    c = parity.Candidate(cuda_lib, n, seed)
    try:
        o.reset(seed, 4, 3, difficulty, max_steps)
        c.reset(seed, 4, 3, difficulty, max_steps)
        merged_dev = torch.zeros(n * layout.action_mask_dtype.itemsize, dtype=torch.uint8, device="cuda")
        dones = wins = 0
        for s in range(steps):
            if regime == "A":
                o.sample()
                c.sample()
            else:
                m = parity.merged_masks(parity.oracle_snapshot(o))
                o.sample(m)
                c.vec.merged_masks_device(merged_dev.data_ptr())
                c.smp.sample_device(merged_dev.data_ptr())
                if s % 50 == 0:
                    c.vec.sync()
                    got = merged_dev.cpu().numpy().view(layout.action_mask_dtype)
                    assert np.array_equal(pb(got), pb(m)), "step %d: device-merged mask" % s
            o.step()
            c.step()
            bad = _small_mismatch(c, o)
            assert bad is None, "step %d: %s" % (s, bad)
            d = np.asarray(o.dones)
            dones += int(d.sum())
            wins += int((np.asarray(o.rewards)[d] != 0).any(axis=1).sum()) if d.any() else 0
            if (s + 1) % 250 == 0 or s == steps - 1:
                bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o))
                assert bad is None, "step %d: %s" % (s, bad)
        if regime == "M":
            assert dones > 100
            if max_steps > 1000:
                assert wins > 100
    finally:
        c.close()
        o.close()


# ---- (c) configs[2] at its real size against the oracle ----------------------------------------
@pytest.mark.parametrize("merged,k", [(False, 200), (True, 60)])
def test_config2_full_size_vs_oracle(cuda_lib, oracle_cls, merged, k):
    """65 536 envs, 4 players, MEDIUM, seed 12345: the fused kernel == k oracle steps on EVERY env."""
    n, seed = 65536, 12345
    o = oracle_cls(n, seed)
    c = parity.Candidate(cuda_lib, n, seed)
    try:
        o.reset(seed, 4, 3, layout.MEDIUM, 100000)
        c.reset(seed, 4, 3, layout.MEDIUM, 100000)
        for _ in range(k):
            if merged:
                o.sample(parity.merged_masks(parity.oracle_snapshot(o)))
            else:
                o.sample()
            o.step()
        c.vec.run_fused(c.smp, 1, merged=merged)
        c.vec.run_fused(c.smp, k - 1, merged=merged)
        bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o))
        assert bad is None, bad
        assert c.vec.stats()["env_steps"] == n * k
    finally:
        c.close()
        o.close()


# ---- (d) configs[4]: the env windows the eight ranks of a 1 M-env job own ----------------------
@pytest.mark.parametrize("rank", range(8))
def test_config4_rank_offsets(cuda_lib, oracle_cls, rank):
    """1 048 576 envs over 8 GPUs = 131 072 per rank; a 4 096-env window at each rank's offset, HARD:
    reset, fused launch (selected masks), then single steps from merged masks (wins, auto-resets)."""
    n, seed = 4096, 12345
    offset = rank * 131072 + (0 if rank % 2 == 0 else 131072 - n)  # first / last window of the rank's slice
    o = oracle_cls(n, seed, offset)
    c = parity.Candidate(cuda_lib, n, seed, offset)
    try:
        o.reset(seed, 4, 3, layout.HARD, 25)
        c.reset(seed, 4, 3, layout.HARD, 25)
        assert parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o)) is None, "after reset"
        for _ in range(150):
            o.sample()
            o.step()
        c.vec.run_fused(c.smp, 150)
        bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o))
        assert bad is None, "fused: " + bad
        for _ in range(100):
            o.sample(parity.merged_masks(parity.oracle_snapshot(o)))
            o.step()
        c.vec.run_fused(c.smp, 100, merged=True)
        bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o))
        assert bad is None, "fused, merged masks: " + bad
        assert c.vec.stats()["episodes"] > 0
    finally:
        c.close()
        o.close()


# ---- (e) boards of 4..8 pieces -------------------------------------------------------------------
@pytest.mark.parametrize("difficulty", [layout.MEDIUM, layout.HARD])
@pytest.mark.parametrize("n_pieces", [4, 5, 6, 7, 8])
def test_boards_many_pieces(cuda_lib, oracle_cls, difficulty, n_pieces):
    """boards of up to 8 travel pieces reach the edge of the 48x48 grid; the retry stack, the
    placed-piece list and the toroidal occupancy bitmap are what they stress.  A batch holding a board that
    leaves the grid must fail on both sides (COG_ERR_MAP_BOUNDS; UB in the reference)."""
    n = 256
    compared = 0
    for seed in (11, 4000, 77000, 123456, 900000, 2000000):
        o = oracle_cls(n, 0)
        c = parity.Candidate(cuda_lib, n, 0)
        try:
            try:
                o.reset(seed, 4, n_pieces, difficulty, 100)
            except port.MapGenerationError:
                with pytest.raises(capi.MapGenerationError):
                    c.reset(seed, 4, n_pieces, difficulty, 100)
                continue
            c.reset(seed, 4, n_pieces, difficulty, 100)
            bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(o))
            assert bad is None, "seed %d: %s" % (seed, bad)
            compared += 1
        finally:
            c.close()
            o.close()
    assert compared >= 4


def test_easy_four_pieces_fails(cuda_lib):
    """EASY has three travel pieces: n_pieces >= 4 always throws (src/tests/test_environment.cpp:68-80)"""
    c = parity.Candidate(cuda_lib, 64, 0)
    for n_pieces in (4, 8):
        with pytest.raises(capi.MapGenerationError):
            c.reset(5, 4, n_pieces, layout.EASY, 100)
    c.close()


# ---- (f) explicit reset() in the middle of an episode -------------------------------------------
@pytest.mark.parametrize("difficulty,max_steps", [(layout.EASY, 100000), (layout.HARD, 40)])
def test_mid_episode_reset(cuda_lib, oracle_cls, difficulty, max_steps):
    """reset() without arguments keeps the RNG streams, does not clear `played`,
    idx_last_activated or n_in_market; then the game goes on."""
    parity.run_mid_episode_reset(lambda n, seed: parity.Candidate(cuda_lib, n, seed), oracle_cls, 256, 4321,
                                 difficulty, max_steps)


# ---- (g) two players, against the reference where it is defined ---------------------------------
def test_two_players_vs_reference(cuda_lib):
    """oracle: the compiled reference (N = 1 instances) on envs whose every board in the window drew start
    piece A — the only 2-player games the reference defines (src/map.cpp:343-354, SURVEY A#13)."""
    if not refshim.available():
        pytest.skip("oracle/_ref not built")
    seeds = parity.two_player_clean_seeds(99000, 8192, layout.HARD, 20, 120, want=12)
    assert len(seeds) >= 6
    for seed in seeds:
        r = refshim.RefVec(1, seed)
        c = parity.Candidate(cuda_lib, 1, seed)
        try:
            r.reset(seed, 2, 3, layout.HARD, 20)
            c.reset(seed, 2, 3, layout.HARD, 20)
            for s in range(120):
                m = r.merged_masks()
                r.sample(m)
                c.sample(m)
                r.step()
                c.step()
                bad = parity.first_mismatch(c.snapshot(), parity.oracle_snapshot(r))
                assert bad is None, "seed %d step %d: %s" % (seed, s, bad)
        finally:
            c.close()
            r.close()
