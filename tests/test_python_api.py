"""The drop-in Python surface (csrc/pybind_module.cpp): same names and call shapes as the
reference's `city_of_gold` module (src/pybind/*.cpp), driven with the reference's own
fuzz loop (pytest/test_vec_env.py:74-113).  CPU: the module linked against the emulated
library (test infrastructure).  GPU: the real module."""
import numpy as np
import pytest

import parity
from gym_eldorado_b200 import layout
from gym_eldorado_b200.layout import payload_bytes as pb


def run_test(cog, steps, n_envs, seed, threaded=False, threads=None, sync=False, oracle=None):
    """pytest/test_vec_env.py:74-113, line for line where the API allows."""
    vec = cog.vec
    env_cls = vec.get_vec_env(n_envs)
    sampler_cls = vec.get_vec_sampler(n_envs)
    envs = env_cls()
    samplers = sampler_cls(seed)
    if threaded:
        runner_cls = vec.get_runner(n_envs)
        runner = runner_cls(envs, samplers, threads)
        step_fun = (lambda _: runner.step_sync()) if sync else (lambda _: runner.step())
        sample_fun = lambda _: runner.sample()
        envs = runner.get_envs()
    else:
        step_fun = envs.step
        sample_fun = samplers.sample
    envs.reset(seed, 4, 3, cog.Difficulty.EASY, 100000, False)
    actions = samplers.get_actions()
    next_obs = envs.observations            # persistent device views, fetched once
    player_masks = envs.selected_action_masks
    current_rewards = envs.rewards
    current_dones = envs.dones
    current_infos = envs.infos
    agents = envs.agent_selection
    for i in range(steps):
        sample_fun(player_masks)
        step_fun(actions)
        if oracle is not None:
            oracle.sample(); oracle.step()
    if threaded:
        runner.sync()
    return dict(observations=next_obs.numpy(), selected_action_masks=player_masks.numpy(),
                rewards=current_rewards.numpy(), dones=current_dones.numpy(), infos=current_infos.numpy(),
                agent_selection=agents.numpy(), actions=actions.numpy())


def _check_api(cog, oracle_cls, steps, n):
    assert cog.Difficulty.EASY == cog.EASY and int(cog.Difficulty.HARD) == 2
    assert cog.vec.get_vec_env(n).__name__ == "vec_cog_env_%d" % n
    assert cog.vec.get_vec_env(n) is cog.vec.get_vec_env(n)
    for mode in ((False, None, False), (True, 4, False), (True, 4, True)):
        o = oracle_cls(n, 123456)
        o.reset(123456, 4, 3, layout.EASY, 100000)
        snap = run_test(cog, steps, n, 123456, *mode, oracle=o)
        assert parity.first_mismatch(snap, parity.oracle_snapshot(o)) is None, mode
        assert snap["observations"].dtype == layout.obs_dtype
        assert snap["rewards"].shape == (n, 4) and snap["dones"].dtype == np.bool_
        o.close()
    # host numpy arrays are accepted where the reference takes ndarrays
    envs = cog.vec.get_vec_env(n)()
    smp = cog.vec.get_vec_sampler(n)(7)
    envs.reset(7, 4, 3, cog.Difficulty.HARD, 50, False)
    o = oracle_cls(n, 7)
    o.reset(7, 4, 3, layout.HARD, 50)
    for _ in range(20):
        m = envs.selected_action_masks.numpy()
        smp.sample(m)
        a = smp.get_actions().numpy()
        envs.step(a)
        o.sample(); o.step()
    assert np.array_equal(pb(envs.observations.numpy()), pb(o.observations))
    # fused runner == the same number of single steps
    r = cog.vec.get_runner(n)(envs, smp)
    r.run(25)
    for _ in range(25):
        o.sample(); o.step()
    r.sync()
    assert np.array_equal(pb(envs.observations.numpy()), pb(o.observations))
    assert envs.episode_stats()["env_steps"] == n * 45
    # merged legal masks (not in the reference): a policy-in-the-loop step and the fused merged run
    for _ in range(30):
        m = envs.merged_action_masks()
        assert np.array_equal(pb(m.numpy()), pb(parity.merged_masks({
            "observations": envs.observations.numpy(), "agent_selection": envs.agent_selection.numpy(),
            "selected_action_masks": envs.selected_action_masks.numpy()})))
        smp.sample(m)                      # device buffer in, zero copy
        envs.step(smp.get_actions())
        o.sample(parity.merged_masks(parity.oracle_snapshot(o))); o.step()
    r.run(40, merged_masks=True)
    for _ in range(40):
        o.sample(parity.merged_masks(parity.oracle_snapshot(o))); o.step()
    r.sync()
    assert np.array_equal(pb(envs.observations.numpy()), pb(o.observations))
    assert np.array_equal(envs.rewards.numpy(), o.rewards)
    # compact observation, one-byte hex grid and the rollout ring (not in the reference)
    assert np.array_equal(layout.grid_to_map(envs.hex_grid.numpy()), o.observations["shared"]["map"])
    for merged in (True, False):
        rec = envs.compact_observations(merged_masks=merged)
        assert rec.numpy().dtype == layout.compact_dtype
        assert parity.compact_mismatch(rec.numpy(), parity.compact_expected(o, merged)) is None
    envs.attach_rollout(8)
    want = np.zeros((n, 8), dtype=layout.compact_dtype)
    for t in range(13):
        rec = parity.compact_expected(o, True)
        o.sample(parity.merged_masks(parity.oracle_snapshot(o)))
        rec["action"] = pb(o.actions)
        o.step()
        rec["done"], rec["filled"], rec["t"] = o.dones, 1, t
        rec["rewards"] = np.where(o.dones[:, None], o.rewards, 0)
        want[:, t % 8] = rec
    r.run(13, merged_masks=True)
    assert envs.rollout_steps == 13
    ring = envs.rollout
    assert ring.numpy().shape == (n, 8) and parity.compact_mismatch(ring.numpy(), want) is None
    if cog.__name__.endswith("_emu"):  # host pointers: numpy can consume the capsule
        assert np.from_dlpack(ring).shape == (n, 8, 256)
    envs.attach_rollout(0)
    with pytest.raises(ValueError):
        envs.rollout
    with pytest.raises(RuntimeError):
        envs.reset(123, 3, 4, cog.Difficulty.EASY, 100, False)  # generate_map_failure
    with pytest.raises(ValueError):
        envs.step(np.zeros(1, dtype=layout.action_dtype))
    o.close()


def test_python_api_emulated(oracle_cls):
    cog = parity.emu_module()
    _check_api(cog, oracle_cls, 150, 16)
    # DLPack export (kDLCPU under the emulator): zero-copy view of the library's buffer
    envs = cog.vec.get_vec_env(4)()
    envs.reset(1, 4, 3, cog.Difficulty.EASY, 10, False)
    v = np.from_dlpack(envs.observations)
    assert v.shape == (4, layout.obs_dtype.itemsize) and v.dtype == np.uint8
    assert np.array_equal(v.view(layout.obs_dtype).reshape(4)["shared"]["shop"], envs.observations.numpy()["shared"]["shop"])
    assert np.from_dlpack(envs.rewards).shape == (4, 4)
    import torch  # the consumer the GPU path is meant for; here it sees a CPU capsule
    t = torch.from_dlpack(envs.observations)
    assert t.shape == (4, layout.obs_dtype.itemsize) and t.dtype == torch.uint8
    assert t.data_ptr() == envs.observations.device_pointer
    assert torch.from_dlpack(envs.dones).dtype == torch.bool


@pytest.mark.gpu
def test_python_api_gpu(cuda_lib, oracle_cls):
    import torch
    from gym_eldorado_b200 import city_of_gold as cog
    _check_api(cog, oracle_cls, 1000, 16)
    n = 256
    envs = cog.vec.get_vec_env(n)()
    envs.reset(5, 4, 3, cog.Difficulty.MEDIUM, 100000, False)
    obs_t = torch.from_dlpack(envs.observations)          # zero-copy device views
    masks_t = torch.as_tensor(envs.selected_action_masks, device="cuda")
    assert obs_t.is_cuda and obs_t.shape == (n, layout.obs_dtype.itemsize) and obs_t.dtype == torch.uint8
    assert obs_t.data_ptr() == envs.observations.device_pointer
    assert masks_t.data_ptr() == envs.selected_action_masks.device_pointer
    host = envs.observations.numpy()
    assert np.array_equal(obs_t.cpu().numpy().view(layout.obs_dtype).reshape(n)["shared"]["shop"], host["shared"]["shop"])
    # a torch CUDA tensor of ActionData records is accepted by step()
    smp = cog.vec.get_vec_sampler(n)(5)
    smp.sample(masks_t)
    acts = torch.from_dlpack(smp.get_actions()).clone()
    envs.step(acts)
    envs.sync()
    assert torch.from_dlpack(envs.rewards).shape == (n, 4)
    assert torch.from_dlpack(envs.dones).dtype == torch.bool
