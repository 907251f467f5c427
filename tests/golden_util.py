"""Replay the golden trajectories (tests/golden/, minted from the unmodified reference by
tools/make_golden.py) on any implementation with reset/sample/step + a snapshot."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_golden  # noqa: E402

GOLD = make_golden.GOLD
TRAJECTORIES = make_golden.TRAJECTORIES


def check_trajectory(name, vec, snapshot, merged, steps=None):
    cfg = list(TRAJECTORIES[name])
    gold = np.load(os.path.join(GOLD, "traj_%s.npz" % name))
    full = steps is None or steps >= cfg[6]
    if not full:
        cfg[6] = steps
    digests, final = make_golden.drive(vec, tuple(cfg), snapshot, merged)
    bad = np.flatnonzero(digests != gold["digests"][:len(digests)])
    assert len(bad) == 0, "%s: first digest mismatch at step %d" % (name, bad[0])
    if full:
        for k, v in final.items():
            assert np.array_equal(v, gold["final_" + k]), "%s: final %s differs" % (name, k)


def check_boards(make_vec, snapshot_of):
    """make_vec(n) -> vec with reset(); snapshot_of(vec) -> dict of buffers."""
    from gym_eldorado_b200.layout import payload_bytes as pb
    gold = np.load(os.path.join(GOLD, "boards.npz"))
    for name, diff in (("EASY", 0), ("MEDIUM", 1), ("HARD", 2)):
        v = make_vec(48)
        v.reset(1000, 4, 3, diff, 100)
        s = snapshot_of(v)
        assert np.array_equal(pb(s["observations"]), gold["obs_" + name]), name
        assert np.array_equal(pb(s["selected_action_masks"]), gold["masks_" + name]), name
        v.close()
