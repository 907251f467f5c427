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
    _check_reference_idioms(cog, oracle_cls, n)
    _check_single_env(cog)


def _check_reference_idioms(cog, oracle_cls, n):
    """What unmodified reference user code does with the views (README / pytest/test_vec_env.py:97-113):
    np.nonzero(dones), infos[i], np.asarray(view); and the host-array call shapes with result arrays."""
    envs = cog.vec.get_vec_env(n)()
    smp = cog.vec.get_vec_sampler(n)(11)
    envs.reset(11, 4, 3, cog.Difficulty.HARD, 6, False)
    o = oracle_cls(n, 11)
    o.reset(11, 4, 3, layout.HARD, 6)
    dones_view, infos_view, masks_view = envs.dones, envs.infos, envs.selected_action_masks
    masks = np.zeros(n, dtype=layout.action_mask_dtype)
    actions = np.zeros(n, dtype=layout.action_dtype)
    rewards = np.zeros((n, 4), dtype=np.float32)
    dones = np.zeros(n, dtype=np.bool_)
    agents = np.zeros(n, dtype=np.uint8)
    masks[...] = np.asarray(masks_view)
    seen = 0
    for s in range(80):
        smp.sample(masks, out=actions)                       # host masks in, host actions out
        o.sample()
        assert np.array_equal(pb(actions), pb(o.actions))
        envs.step(actions, out_masks=masks, out_rewards=rewards, out_dones=dones, out_agents=agents)
        actions[...] = 0                                     # the input is free again as soon as step() returns
        o.step()
        assert np.array_equal(pb(masks), pb(o.selected_action_masks)) and np.array_equal(dones, o.dones)
        assert np.array_equal(rewards, o.rewards) and np.array_equal(agents, o.agent_selection)
        for i in np.nonzero(dones_view)[0]:                  # the reference's idiom, on the device view
            assert infos_view[i]["total_length"] == o.infos[i]["total_length"] == 6
            seen += 1
        assert np.array_equal(np.asarray(dones_view), o.dones)
    assert seen > 0 and len(dones_view) == n
    assert np.asarray(envs.rewards).shape == (n, 4) and np.asarray(envs.observations).dtype == layout.obs_dtype
    o.close()


def _check_single_env(cog):
    """city_of_gold.cog_env / action_sampler (reference src/pybind/single_env.cpp:13-32, common.cpp:25-27):
    N = 1 handles with the single env's semantics (no auto-reset, steps on a finished env are ignored),
    against the reference's own cog_env + action_sampler where oracle/_ref is built, else the C port."""
    from oracle import port, refshim
    seed, n_players, diff, max_steps = 7, 4, layout.HARD, 12
    env = cog.cog_env(seed, n_players, 3, cog.Difficulty.HARD, max_steps, False)
    smp = cog.action_sampler(seed)
    assert env.get_seed() == seed and env.get_n_players() == n_players
    env.reset()
    if refshim.available():
        ref = refshim.RefSingle(seed, n_players, 3, diff, max_steps)
        ref.reset()
        ref_smp = refshim.RefSampler(seed)
        ref_done = lambda: ref.done
        ref_agent = lambda: ref.agent_selection
    else:
        ref = port.OracleVec(1, seed)
        ref.set_autoreset(False)
        ref.reset(seed, n_players, 3, diff, max_steps)
        ref_smp = None
        ref_done = lambda: bool(ref.dones[0])
        ref_agent = lambda: int(ref.agent_selection[0])
    steps = 0
    for episode in range(3):
        while not env.get_done() and steps < 2000:
            obs = env.observations
            agent = env.agent_selection
            # merged legal mask of the current agent, assembled from the record classes
            m = np.zeros(1, dtype=layout.action_mask_dtype)
            S, P = env.selected_action_mask, obs.player_data[agent]
            m["play"], m["play_special"] = S.play, S.play_special
            m["move"], m["get_from_shop"] = P.action_mask.move, P.action_mask.get_from_shop
            rem = np.array(P.action_mask.remove)
            rem[1:] &= np.asarray(P.obs.hand) > 0
            m["remove"] = rem
            a = smp.sample(cog.ActionMask(m[0]))
            rec = np.zeros(1, dtype=layout.action_dtype)
            for f in layout.action_dtype.names:
                rec[f] = getattr(a, f)
            if ref_smp is not None:
                assert np.array_equal(pb(ref_smp.sample(m)), pb(rec)), steps
            else:
                ref.sample(m)
                assert np.array_equal(pb(ref.actions), pb(rec)), steps
            env.step(a)
            ref.step(rec)
            steps += 1
            assert env.get_done() == ref_done() and env.agent_selection == ref_agent(), steps
            assert np.array_equal(pb(np.asarray(env.observations.shared.map)), pb(ref.observations["shared"]["map"][0]))
            assert np.array_equal(np.asarray(env.observations.shared.resources), ref.observations["shared"]["current_resources"][0])
            assert np.array_equal(np.asarray(env.observations.player_data[agent].obs.hand),
                                  ref.observations["player_data"][0][agent]["obs"]["hand"])
            assert np.array_equal(np.asarray(env.selected_action_mask.play), ref.selected_action_masks["play"][0])
        assert env.get_done()
        assert np.array_equal(np.asarray(env.rewards), ref.rewards[0])
        assert int(env.info.total_length) == int(ref.infos["total_length"][0]) == max_steps
        before = np.asarray(env.observations.player_data[0].obs.discard).copy()
        env.step(cog.ActionData(play=1))       # dead_step: ignored (src/environment.cpp:92-95)
        assert env.get_done() and np.array_equal(before, np.asarray(env.observations.player_data[0].obs.discard))
        env.reset()
        ref.reset()
        assert not env.get_done()
        assert np.array_equal(np.asarray(env.observations.shared.map), ref.observations["shared"]["map"][0])
    assert steps > 100
    # reset(seed, ...) reseeds; default-constructed envs work too
    env.reset(99, 4, 3, cog.Difficulty.EASY, 50, False)
    assert env.get_seed() == 99 and env.agent_selection == 0
    cog.cog_env().reset()
    if hasattr(ref, "close"):
        ref.close()


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
    # a pinned host array is free again as soon as step() / sample() return (the copy is really asynchronous
    # there; the reference's calls are synchronous with respect to their input)
    n2 = 4096
    envs2 = cog.vec.get_vec_env(n2)()
    smp2 = cog.vec.get_vec_sampler(n2)(77)
    envs2.reset(77, 4, 3, cog.Difficulty.HARD, 9, False)
    o = oracle_cls(n2, 77)
    o.reset(77, 4, 3, layout.HARD, 9)
    pin_a = torch.zeros(n2 * layout.action_dtype.itemsize, dtype=torch.uint8).pin_memory()
    pin_m = torch.zeros(n2 * layout.action_mask_dtype.itemsize, dtype=torch.uint8).pin_memory()
    a_host, m_host = pin_a.numpy().view(layout.action_dtype), pin_m.numpy().view(layout.action_mask_dtype)
    for _ in range(40):
        m_host[...] = envs2.selected_action_masks.numpy()
        smp2.sample(m_host)
        m_host.view(np.uint8)[...] = 1                      # garbage right after the call
        a_host[...] = smp2.get_actions().numpy()
        envs2.step(a_host)
        a_host.view(np.uint8)[...] = 255
        o.sample(); o.step()
    assert parity.first_mismatch(
        dict(observations=envs2.observations.numpy(), selected_action_masks=envs2.selected_action_masks.numpy(),
             rewards=envs2.rewards.numpy(), dones=envs2.dones.numpy(), infos=envs2.infos.numpy(),
             agent_selection=envs2.agent_selection.numpy(), actions=smp2.get_actions().numpy()),
        parity.oracle_snapshot(o)) is None
    o.close()
    # a consumer on its own stream: __dlpack__(stream=...) orders it after the library's work
    side = torch.cuda.Stream()
    envs2.step(smp2.get_actions())
    with torch.cuda.stream(side):
        d = torch.from_dlpack(envs2.dones).clone()
    side.synchronize()
    assert np.array_equal(d.cpu().numpy(), envs2.dones.numpy())
    # the packed host loop through the ctypes layer is covered in test_gpu_parity.py::test_packed_host_io
