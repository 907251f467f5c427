"""Shared parity driver: the C-ABI library (CUDA on a GPU box, or the emulated build
in tests/emu for CPU-only logic checks) against the oracle, payload bytes only."""
import os
import subprocess

import numpy as np

from gym_eldorado_b200 import capi, layout
from gym_eldorado_b200.layout import payload_bytes as pb

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_PATH = os.path.join(HERE, "emu", "libcog_emu.so")
_emu = None


def emu_lib():
    """TEST-ONLY: the kernels compiled for the host under the lock-step warp emulator."""
    global _emu
    if _emu is None:
        subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "emu")])
        _emu = capi.load(EMU_PATH)
    return _emu


class Candidate:
    """vec env + sampler driven through the C ABI, with host snapshots of every buffer."""

    def __init__(self, library, n, sampler_seed, offset=0):
        self.n = n
        self.vec = capi.VecHandle(n, offset, 0, None, library)
        self.smp = capi.SamplerHandle(n, sampler_seed, offset, 0, None, library)
        self.masks_ptr = self.vec.buffer(layout.BUF_SELECTED_MASKS)[0]

    def reset(self, *a, **k):
        self.vec.reset(*a, **k)

    def sample(self, masks=None):
        if masks is None:
            self.smp.sample_device(self.masks_ptr)
        else:
            self.smp.sample_host(np.ascontiguousarray(masks))

    def step(self):
        self.vec.step_device(self.smp.actions_ptr())

    def step_sampled(self):
        """sample(selected masks); step — one kernel (cog_vec_step_sampled)"""
        self.vec.step_sampled(self.smp)

    def snapshot(self):
        v = self.vec
        return {
            "observations": v.read(layout.BUF_OBS),
            "selected_action_masks": v.read(layout.BUF_SELECTED_MASKS),
            "rewards": v.read(layout.BUF_REWARDS),
            "dones": v.read(layout.BUF_DONES),
            "infos": v.read(layout.BUF_INFOS),
            "agent_selection": v.read(layout.BUF_AGENT_SELECTION),
            "actions": self.smp.read_actions(),
        }

    def merged(self):
        """cog_vec_merged_masks into a scratch buffer (device memory on the GPU box, host memory
        under the emulator), returned as a host ActionMask array."""
        nbytes = self.n * layout.action_mask_dtype.itemsize
        if self.vec.L is _emu:
            out = np.zeros(self.n, dtype=layout.action_mask_dtype)
            self.vec.merged_masks_device(out.ctypes.data)
            return out
        import torch
        t = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
        self.vec.merged_masks_device(t.data_ptr())
        self.vec.stats()  # synchronises the handle's stream
        return t.cpu().numpy().view(layout.action_mask_dtype)

    def close(self):
        self.vec.close()
        self.smp.close()


def oracle_snapshot(o):
    return {k: getattr(o, k) for k in ("observations", "selected_action_masks", "rewards", "dones",
                                       "infos", "agent_selection", "actions")}


def merged_masks(snap):
    """Regime M: S.play/play_special, P[agent].move/get_from_shop,
    P[agent].remove AND hand > 0 — exercises every branch of step."""
    n = len(snap["agent_selection"])
    idx = np.arange(n)
    pd = snap["observations"]["player_data"][idx, snap["agent_selection"]]
    m = np.zeros(n, dtype=layout.action_mask_dtype)
    S = snap["selected_action_masks"]
    m["play"] = S["play"]
    m["play_special"] = S["play_special"]
    m["move"] = pd["action_mask"]["move"]
    m["get_from_shop"] = pd["action_mask"]["get_from_shop"]
    rem = pd["action_mask"]["remove"].copy()
    rem[:, 1:] &= pd["obs"]["hand"] > 0
    m["remove"] = rem
    return m


def player_masks(snap):
    """Regime P: the current agent's own ActionMask record, as the reference's C++ doctests
    use it — legal-looking plays the shared mask would have refused (u8 pile counters wrap)."""
    n = len(snap["observations"])
    agent = np.asarray(snap["agent_selection"]).astype(np.int64)
    return snap["observations"]["player_data"][np.arange(n), agent]["action_mask"].copy()


def first_mismatch(a, b):
    for name in ("observations", "selected_action_masks", "infos", "actions"):
        x, y = pb(a[name]), pb(b[name])
        if not np.array_equal(x, y):
            bad = np.argwhere(x != y)
            return "%s: env %d payload byte %d: %d != %d (%d bytes differ)" % (
                name, bad[0][0], bad[0][1], x[tuple(bad[0])], y[tuple(bad[0])], len(bad))
    for name in ("rewards", "dones", "agent_selection"):
        if not np.array_equal(np.asarray(a[name]), np.asarray(b[name])):
            return name
    return None


def run_parity(library, oracle_cls, n, seed, difficulty, max_steps, steps, regime, n_players=4,
               n_pieces=3, check_every=1, offset=0):
    """Drive candidate and oracle side by side; returns the number of dones seen.
    regime "S" = regime A through the one-kernel cog_vec_step_sampled."""
    o = oracle_cls(n, seed) if offset == 0 else oracle_cls(n, seed, offset)
    c = Candidate(library, n, seed, offset)
    try:
        o.reset(seed, n_players, n_pieces, difficulty, max_steps)
        c.reset(seed, n_players, n_pieces, difficulty, max_steps)
        snap = c.snapshot()
        bad = first_mismatch(snap, oracle_snapshot(o))
        assert bad is None, "after reset: " + bad
        dones = 0
        for s in range(steps):
            if regime == "A":
                o.sample()
                c.sample()
            elif regime == "S":
                o.sample()
            else:
                if check_every != 1:
                    snap = c.snapshot()
                m = player_masks(snap) if regime == "P" else merged_masks(snap)
                o.sample(m)
                c.sample(m)
            o.step()
            if regime == "S":
                c.step_sampled()
            else:
                c.step()
            dones += int(o.dones.sum())
            if regime not in ("A", "S") or (s + 1) % check_every == 0 or s == steps - 1:
                snap = c.snapshot()
            if (s + 1) % check_every == 0 or s == steps - 1:
                bad = first_mismatch(snap, oracle_snapshot(o))
                assert bad is None, "step %d: %s" % (s, bad)
        return dones
    finally:
        c.close()
        o.close()


def emu_module():
    """TEST-ONLY: the pybind host module (csrc/pybind_module.cpp) linked against the
    emulated library, so the reference-style Python loop can be exercised without a GPU."""
    import importlib.util
    import sysconfig
    from gym_eldorado_b200 import pybind_build
    emu_lib()
    emu_dir = os.path.join(HERE, "emu")
    target = os.path.join(emu_dir, "city_of_gold_emu" + sysconfig.get_config_var("EXT_SUFFIX"))
    deps = [pybind_build.SRC, EMU_PATH]
    if not os.path.exists(target) or any(os.path.getmtime(d) > os.path.getmtime(target) for d in deps):
        pybind_build.compile_module(target, emu_dir, "libcog_emu.so", "city_of_gold_emu",
                                    extra=("-DCOG_HOST_POINTERS",))
    spec = importlib.util.spec_from_file_location("city_of_gold_emu", target)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run_fused_merged_parity(library, oracle_cls, n, seed, difficulty, max_steps, chunks, n_players=4):
    """Fused kernel sampling from the merged legal mask (regime M) against the oracle driven
    step by step with host-merged masks; also checks cog_vec_merged_masks itself."""
    o = oracle_cls(n, seed)
    c = Candidate(library, n, seed)
    try:
        o.reset(seed, n_players, 3, difficulty, max_steps)
        c.reset(seed, n_players, 3, difficulty, max_steps)
        for k in chunks:
            for _ in range(k):
                o.sample(merged_masks(oracle_snapshot(o)))
                o.step()
            c.vec.run_fused(c.smp, k, merged=True)
            snap = c.snapshot()
            bad = first_mismatch(snap, oracle_snapshot(o))
            assert bad is None, bad
            want, got = merged_masks(snap), c.merged()
            assert np.array_equal(pb(want), pb(got)), "cog_vec_merged_masks differs from the host merge"
        return c.vec.stats()
    finally:
        c.close()
        o.close()


def _host_array(shape, dtype, pinned):
    """Zeroed host array; pinned=True: page-locked memory the kernels can address (torch's
    pin_memory()), which takes the zero-copy branch of the *_host_io entry points."""
    if not pinned:
        return np.zeros(shape, dtype=dtype)
    import torch
    dt = np.dtype(dtype)
    count = int(np.prod(shape))
    t = torch.zeros(count * dt.itemsize, dtype=torch.uint8).pin_memory()
    return t.numpy().view(dt).reshape(shape)  # the view keeps the pinned tensor alive through .base


def run_host_io_parity(library, oracle_cls, n, seed, difficulty, max_steps, steps, regime="A", pinned=False):
    """The host-buffer entry points (cog_sampler_sample_host_io, cog_vec_step_host_io) driven like
    the reference's Python loop — masks and actions only ever live in host arrays — against the
    oracle.  Pageable arrays go through the chunked copy/compute pipeline, pinned ones are read
    and written by the kernels directly."""
    o = oracle_cls(n, seed)
    c = Candidate(library, n, seed)
    try:
        o.reset(seed, 4, 3, difficulty, max_steps)
        c.reset(seed, 4, 3, difficulty, max_steps)
        masks = _host_array(n, layout.action_mask_dtype, pinned)
        c.vec.read(layout.BUF_SELECTED_MASKS, out=masks)
        actions = _host_array(n, layout.action_dtype, pinned)
        rewards = _host_array((n, 4), np.float32, pinned)
        dones = _host_array(n, np.bool_, pinned)
        agents = _host_array(n, np.uint8, pinned)
        merged = _host_array(n, layout.action_mask_dtype, pinned)
        n_done = 0
        for s in range(steps):
            if regime == "M":
                merged[...] = merged_masks(c.snapshot())
                m = merged
                o.sample(m)
            else:
                m = masks
                o.sample()
            c.smp.sample_host_io(m, actions)
            assert np.array_equal(pb(actions), pb(o.actions)), "step %d: actions" % s
            o.step()
            c.vec.step_host_io(actions, masks, rewards, dones, agents)
            assert np.array_equal(pb(masks), pb(o.selected_action_masks)), "step %d: masks" % s
            assert np.array_equal(rewards, o.rewards) and np.array_equal(dones, o.dones), "step %d" % s
            assert np.array_equal(agents, o.agent_selection), "step %d: agent_selection" % s
            n_done += int(dones.sum())
        bad = first_mismatch(c.snapshot(), oracle_snapshot(o))
        assert bad is None, bad
        return n_done
    finally:
        c.close()
        o.close()


def _bits(a):
    a = np.asarray(a).astype(np.uint64)
    return (a << np.arange(a.shape[1], dtype=np.uint64)).sum(axis=1).astype(np.uint32)


def compact_expected(o, merged, n_players=4):
    """The compact record (include/cog_b200.h, cog_vec_compact_obs) of every env's current
    state, restated from the oracle's reference-layout buffers; player positions and the
    turn counter (which the reference never exports) from the oracle's hidden state."""
    snap = oracle_snapshot(o)
    n = o.n
    idx = np.arange(n)
    ag = snap["agent_selection"]
    obs = snap["observations"]
    r = np.zeros(n, dtype=layout.compact_dtype)
    r["agent_selection"] = ag
    r["phase"] = obs["shared"]["phase"]
    r["resources_x2"] = np.round(obs["shared"]["current_resources"] * 2).astype(np.uint16)
    pd = obs["player_data"][idx, ag]
    for name in ("draw", "hand", "active", "played", "discard"):
        r["cards"][name] = pd["obs"][name]
    r["cards"]["shop"][:, :layout.N_BUYABLETYPES] = obs["shared"]["shop"]
    m = merged_masks(snap) if merged else snap["selected_action_masks"]
    r["mask_play"] = _bits(m["play"])
    r["mask_play_special"] = _bits(m["play_special"])
    r["mask_remove"] = _bits(m["remove"])
    r["mask_move_shop"] = _bits(m["move"]) | (_bits(m["get_from_shop"]) << np.uint32(7))
    h = np.stack([o.hidden(i) for i in range(n)]).astype(np.uint32).view(np.int32).astype(np.int64)
    r["turn_counter"] = h[:, 1]
    for p in range(min(n_players, 4)):
        r["loc"]["row"][:, p] = (h[:, 4 + 2 * p] - h[:, 12]) // 2 + 1
        r["loc"]["col"][:, p] = (h[:, 5 + 2 * p] - h[:, 13]) // 2 + 1
    return r


def compact_mismatch(got, want):
    a = np.ascontiguousarray(got).view(np.uint8).reshape(-1, layout.compact_dtype.itemsize)
    b = np.ascontiguousarray(want).view(np.uint8).reshape(-1, layout.compact_dtype.itemsize)
    if np.array_equal(a, b):
        return None
    bad = np.argwhere(a != b)
    return "record %d byte %d: %d != %d (%d bytes differ)" % (bad[0][0], bad[0][1], a[tuple(bad[0])],
                                                               b[tuple(bad[0])], len(bad))


def run_rollout_parity(library, oracle_cls, n, seed, difficulty, max_steps, chunks, n_slots, merged,
                       n_players=4):
    """The rollout ring (cog_vec_rollout_attach + fused launches) and cog_vec_compact_obs
    against records restated from the oracle stepped one call at a time; also the one-byte
    hex grid (BUF_GRID) against the reference-layout map observation."""
    o = oracle_cls(n, seed)
    c = Candidate(library, n, seed)
    try:
        o.reset(seed, n_players, 3, difficulty, max_steps)
        c.reset(seed, n_players, 3, difficulty, max_steps)
        c.vec.rollout_attach(n_slots)
        want = np.zeros((n, n_slots), dtype=layout.compact_dtype)
        t = 0
        dones = wins = 0
        for k in chunks:
            for _ in range(k):
                r = compact_expected(o, merged, n_players)
                if merged:
                    o.sample(merged_masks(oracle_snapshot(o)))
                else:
                    o.sample()
                r["action"] = pb(o.actions)
                o.step()
                r["done"] = o.dones
                r["filled"] = 1
                r["rewards"] = np.where(o.dones[:, None], o.rewards, 0)
                r["t"] = t
                want[:, t % n_slots] = r
                t += 1
                dones += int(o.dones.sum())
                wins += int((r["rewards"] != 0).any(axis=1).sum())
            c.vec.run_fused(c.smp, k, merged=merged)
            bad = first_mismatch(c.snapshot(), oracle_snapshot(o))
            assert bad is None, bad
            _, slots, recorded = c.vec.rollout_info()
            assert slots == n_slots and recorded == t
            bad = compact_mismatch(c.vec.read_rollout(), want)
            assert bad is None, "ring after %d steps: %s" % (t, bad)
            bad = compact_mismatch(c.vec.compact_obs(merged=merged), compact_expected(o, merged, n_players))
            assert bad is None, "compact_obs after %d steps: %s" % (t, bad)
            grid = c.vec.read(layout.BUF_GRID)
            assert np.array_equal(layout.grid_to_map(grid), o.observations["shared"]["map"]), "hex grid"
        c.vec.rollout_attach(0)
        assert c.vec.rollout_info() == (None, 0, 0)
        return dones, wins
    finally:
        c.close()
        o.close()


# ---- the compiled reference as a second oracle ---------------------------------------------
def ref_oracle_cls():
    """oracle/_ref/libcog_ref.so (the UNMODIFIED reference sources, oracle/Makefile) behind the
    OracleVec surface, so that the GPU tests meet the reference's own outputs directly.  Cases the
    reference cannot run are skipped, not faked: N outside the shim's template instantiations, fewer
    than 4 players (heap overflow in Map::add_players, SURVEY A#13), boards of 6+ pieces (may write
    outside the 48x48 observation)."""
    import pytest
    from oracle import port, refshim
    if not refshim.available():
        pytest.skip("oracle/_ref/libcog_ref.so not built (needs /root/reference at build time)")

    class RefOracle(refshim.RefVec):
        def __init__(self, n, sampler_seed, offset=0):
            if n not in refshim.SUPPORTED_N:
                pytest.skip("the reference shim instantiates N in %s only" % (refshim.SUPPORTED_N,))
            base = (int(sampler_seed) & 0xFFFFFFFF) + offset
            if base > 0xFFFFFFFF:
                pytest.skip("sampler seed + offset does not fit the reference's 32-bit constructor argument")
            self._offset = offset
            super().__init__(n, base)

        def reset(self, seed=None, n_players=4, n_pieces=3, difficulty=0, max_steps=100000):
            if seed is not None:
                if n_players < 4:
                    pytest.skip("fewer than 4 players is undefined behaviour in the reference")
                if n_pieces > 5:
                    pytest.skip("boards of 6+ pieces can leave the reference's 48x48 observation (UB)")
                seed = int(seed) + self._offset
            try:
                super().reset(seed, n_players, n_pieces, difficulty, max_steps)
            except RuntimeError as e:
                if "generate map" in str(e):
                    raise port.MapGenerationError(str(e))
                raise

    return RefOracle


def run_mid_episode_reset(make_candidate, oracle_cls, n, seed, difficulty, max_steps, before=70, after=120):
    """>= 50 regime-M steps, reset() WITHOUT arguments in the middle of the episodes (SURVEY App. A#6/A#7:
    `played`, idx_last_activated and n_in_market survive, the RNG streams continue), then on."""
    o = oracle_cls(n, seed)
    c = make_candidate(n, seed)
    snap_c = c.snapshot if hasattr(c, "snapshot") else (lambda: oracle_snapshot(c))
    try:
        o.reset(seed, 4, 3, difficulty, max_steps)
        c.reset(seed, 4, 3, difficulty, max_steps)
        for phase, steps in (("before", before), ("after", after)):
            for s in range(steps):
                m = merged_masks(oracle_snapshot(o))
                o.sample(m)
                c.sample(m)
                o.step()
                c.step()
                if s % 10 == 9 or s == steps - 1:
                    bad = first_mismatch(snap_c(), oracle_snapshot(o))
                    assert bad is None, "%s the reset, step %d: %s" % (phase, s, bad)
            if phase == "before":
                snap = oracle_snapshot(o)
                # the reset really interrupts games: cards in play, purchases made
                assert snap["observations"]["player_data"]["obs"]["active"].sum() > 0
                o.reset()
                c.reset()
                bad = first_mismatch(snap_c(), oracle_snapshot(o))
                assert bad is None, "right after reset(): " + bad
    finally:
        c.close()
        o.close()


def two_player_clean_seeds(seed, n, difficulty, max_steps, steps, want=12):
    """Seeds s = seed + i of 2-player envs whose every board within `steps` regime-M steps drew start piece A,
    screened with the C restatement (whose 2-player rule is defined for both start pieces)."""
    from oracle import port
    o = port.OracleVec(n, seed)
    try:
        o.reset(seed, 2, 3, difficulty, max_steps)
        for _ in range(steps):
            o.sample(merged_masks(oracle_snapshot(o)))
            o.step()
        clean = [i for i in range(n) if o.hidden(i)[14] == 0]
    finally:
        o.close()
    return [seed + i for i in clean[:want]]


def run_packed_io_parity(library, oracle_cls, n, seed, difficulty, max_steps, steps, regime="A", pinned=False):
    """The packed host-I/O entry points (16-byte masks, 8-byte actions) driven like a host policy loop,
    against the oracle at every step; also cog_vec_packed_masks (selected and merged) against the
    packed form of the oracle's records."""
    o = oracle_cls(n, seed)
    c = Candidate(library, n, seed)
    try:
        o.reset(seed, 4, 3, difficulty, max_steps)
        c.reset(seed, 4, 3, difficulty, max_steps)
        pmasks = _host_array(n, layout.packed_mask_dtype, pinned)
        pactions = _host_array(n, layout.packed_action_dtype, pinned)
        rewards = _host_array((n, 4), np.float32, pinned)
        dones = _host_array(n, np.bool_, pinned)
        agents = _host_array(n, np.uint8, pinned)
        c.vec.packed_masks(out=pmasks)
        assert np.array_equal(pmasks, layout.pack_masks(o.selected_action_masks)), "packed masks after reset"
        assert np.array_equal(pb(layout.unpack_masks(pmasks)), pb(o.selected_action_masks))
        n_done = 0
        for s in range(steps):
            if regime == "M":
                m = merged_masks(oracle_snapshot(o))
                o.sample(m)
                pm = c.vec.packed_masks(merged=True)
                assert np.array_equal(pm, layout.pack_masks(m)), "step %d: packed merged masks" % s
            else:
                o.sample()
                pm = pmasks
            c.smp.sample_packed_host_io(pm, pactions)
            want = pb(o.actions)
            got = np.ascontiguousarray(pactions).view(np.uint8).reshape(n, 8)
            assert np.array_equal(got[:, :5], want) and not got[:, 5:].any(), "step %d: actions" % s
            assert np.array_equal(pb(c.smp.read_actions()), want), "step %d: sampler's own ActionData buffer" % s
            o.step()
            c.vec.step_packed_host_io(pactions, pmasks, rewards, dones, agents)
            assert np.array_equal(pmasks, layout.pack_masks(o.selected_action_masks)), "step %d: masks" % s
            assert np.array_equal(rewards, o.rewards) and np.array_equal(dones, o.dones), "step %d" % s
            assert np.array_equal(agents, o.agent_selection), "step %d: agent_selection" % s
            n_done += int(dones.sum())
        bad = first_mismatch(c.snapshot(), oracle_snapshot(o))
        assert bad is None, bad
        assert c.vec.stats()["env_steps"] == n * steps
        return n_done
    finally:
        c.close()
        o.close()
