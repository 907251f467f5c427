#!/usr/bin/env python
"""A small pass over every kernel of the library, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_workload.py
Short episodes (board generation in every kernel that has it), merged masks, rollout ring, host I/O paths."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gym_eldorado_b200 import capi, layout
L = capi.lib()
n = int(os.environ.get("VB_ENVS", "512"))
for diff, max_steps, players in ((layout.EASY, 6, 4), (layout.HARD, 3, 2), (layout.MEDIUM, 0, 3)):
    v = capi.VecHandle(n, 0, 0, None, L); s = capi.SamplerHandle(n, 7, 0, 0, None, L)
    v.reset(11, players, 3, diff, max_steps)
    mp, ap = v.buffer(layout.BUF_SELECTED_MASKS)[0], s.actions_ptr()
    for _ in range(30):
        s.sample_device(mp); v.step_device(ap)
    for _ in range(30):
        v.step_sampled(s)
    v.run_fused(s, 40, False); v.run_fused(s, 40, True)
    v.rollout_attach(8); v.run_fused(s, 20, False); v.run_fused(s, 20, True)
    v.merged_masks_device(None); v.read(layout.BUF_MERGED_MASKS); v.compact_obs(True)
    pm = v.packed_masks(merged=True); pa = np.zeros(n, dtype=layout.packed_action_dtype)
    rew = np.zeros((n, 4), np.float32); done = np.zeros(n, np.bool_); agent = np.zeros(n, np.uint8)
    for _ in range(5):
        s.sample_packed_host_io(pm, pa); v.step_packed_host_io(pa, pm, rew, done, agent)
        s.sample_packed_noalloc(pm, pa); v.step_packed_host_io(pa, pm, rew, done, agent)
    masks = np.zeros(n, dtype=layout.action_mask_dtype); acts = np.zeros(n, dtype=layout.action_dtype)
    v.read(layout.BUF_SELECTED_MASKS, masks)
    for _ in range(3):
        s.sample_host_io(masks, acts); v.step_host_io(acts, masks, rew, done, agent)
    v.reset(); v.sync()
    print("difficulty", diff, "ok", v.stats()["env_steps"], "env-steps,", v.stats()["resets"], "boards")
    v.close(); s.close()
