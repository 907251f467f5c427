#!/usr/bin/env python
"""Mint golden vectors from the UNMODIFIED reference (oracle/_ref/libcog_ref.so, built
from /root/reference by oracle/Makefile) into tests/golden/.  Run in the build
container only (the reference sources do not exist on the GPU box):

    make -C oracle ref && python tools/make_golden.py

Fixtures (payload bytes only — record padding is indeterminate in the reference):
  boards.npz        reset(seed=1000, 4 players, n_pieces 3) on 48 envs per difficulty:
                    full ObsData / selected-mask payload after reset
  traj_<name>.npz   a trajectory driven sequentially (envs.step, not the racy runner):
                    per-step BLAKE2b-64 digest over every exported buffer + the sampled
                    actions, and the full payload snapshot after the last step
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from gym_eldorado_b200 import layout  # noqa: E402
from gym_eldorado_b200.layout import payload_bytes as pb  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# name: (n_envs, seed, n_players, n_pieces, difficulty, max_steps, steps, regime)
TRAJECTORIES = {
    "readme_A": (16, 0, 4, 3, layout.EASY, 100000, 1000, "A"),       # BASELINE configs[0]
    "asv_A": (64, 12345, 4, 3, layout.EASY, 100000, 600, "A"),       # configs[1] flow
    "easy_M": (16, 777, 4, 3, layout.EASY, 100000, 1500, "M"),
    "hard_M": (16, 777, 4, 3, layout.HARD, 60, 1500, "M"),
    "medium_short_M": (16, 12345, 4, 3, layout.MEDIUM, 30, 1000, "M"),
    "hard_reset_heavy_M": (16, 4242, 4, 3, layout.HARD, 8, 800, "M"),  # configs[3] twin, 4 players
}
BUFS = ("observations", "selected_action_masks", "rewards", "dones", "infos", "agent_selection", "actions")


def snap_payload(s):
    out = {}
    for k in BUFS:
        a = np.asarray(s[k])
        out[k] = pb(a) if a.dtype.names else np.ascontiguousarray(a).view(np.uint8).reshape(len(a), -1)
    return out


def digest(s):
    h = hashlib.blake2b(digest_size=8)
    p = snap_payload(s)
    for k in BUFS:
        h.update(p[k].tobytes())
    return int.from_bytes(h.digest(), "little")


def drive(vec, cfg, snapshot, merged):
    """vec: RefVec / OracleVec / parity.Candidate-like (reset, sample, step)."""
    n, seed, n_players, n_pieces, diff, max_steps, steps, regime = cfg
    vec.reset(seed, n_players, n_pieces, diff, max_steps)
    digests = np.zeros(steps, dtype=np.uint64)
    for s in range(steps):
        if regime == "A":
            vec.sample()
        else:
            vec.sample(merged(snapshot()))
        vec.step()
        digests[s] = digest(snapshot())
    return digests, snap_payload(snapshot())


def main():
    from oracle import refshim
    import parity
    os.makedirs(GOLD, exist_ok=True)
    boards = {}
    for name, diff in (("EASY", 0), ("MEDIUM", 1), ("HARD", 2)):
        r = refshim.RefVec(64, 0)  # the shim instantiates N in {1,16,64,...}
        r.reset(1000, 4, 3, diff, 100)
        boards["obs_" + name] = pb(r.observations)[:48]
        boards["masks_" + name] = pb(r.selected_action_masks)[:48]
        r.close()
    np.savez_compressed(os.path.join(GOLD, "boards.npz"), **boards)
    for name, cfg in TRAJECTORIES.items():
        r = refshim.RefVec(cfg[0], cfg[1])
        d, final = drive(r, cfg, lambda: parity.oracle_snapshot(r), parity.merged_masks)
        np.savez_compressed(os.path.join(GOLD, "traj_%s.npz" % name), digests=d,
                            **{"final_" + k: v for k, v in final.items()})
        print(name, "dones in final step:", int(final["dones"].sum()), "digest[-1] = %016x" % d[-1])
        r.close()


if __name__ == "__main__":
    main()
