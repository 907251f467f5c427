"""The measurements behind bench.py's JSON line (kept apart so that bench.py reads as the contract).

Nothing here is a product path: the kernels are called through the C ABI (gym_eldorado_b200.capi), the
reference through oracle/_ref (bench.py --impl reference and the cpu_baseline / cpu_legs only)."""
import json
import os
import subprocess
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

METRIC = "env-steps/sec (random masked policy)"
DTYPE = "u8/u32 integer (f32 exports)"
# payload bytes that must move per sample+step when every exported buffer is
# kept current each step (391 read + 305 written)
ALGO_BYTES_PER_ENV_STEP = 696
ALGO_BYTES_PER_RESET = 19440
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md
N_SMS, SMSP_PER_SM = 148, 4

# BASELINE.json configs (0-based index); [2] is the headline single-GPU workload
WORKLOADS = {
    2: {"workload": "BASELINE configs[2]: 65,536 envs per GPU, 4 players, 3 pieces, MEDIUM, "
                    "max_steps 100000, masked random sampler, persistent fused step+sample kernel",
        "n_envs_per_gpu": 65536, "n_players": 4, "n_pieces": 3, "difficulty": "MEDIUM",
        "max_steps": 100000, "seed": 12345},
    3: {"workload": "BASELINE configs[3] (reset-heavy): 65,536 envs per GPU, 2 players, 3 pieces, HARD, "
                    "max_steps 8 turns, fused kernel with procedural board generation on every episode end",
        "n_envs_per_gpu": 65536, "n_players": 2, "n_pieces": 3, "difficulty": "HARD",
        "max_steps": 8, "seed": 12345},
    4: {"workload": "BASELINE configs[4]: 1,048,576 envs sharded over the GPUs, 4 players, 3 pieces, HARD, "
                    "max_steps 100000, fused kernel, NCCL-reduced episode statistics",
        "n_envs_per_gpu": None, "n_envs_total": 1 << 20, "n_players": 4, "n_pieces": 3, "difficulty": "HARD",
        "max_steps": 100000, "seed": 12345},
}
DIFFICULTY = {"EASY": 0, "MEDIUM": 1, "HARD": 2}


def config_of(args, world):
    """The workload description — the SAME keys and values in both arms (b200 / reference)."""
    cfg = dict(args.workload)
    n = args.envs
    cfg.update({"n_envs_per_gpu": n, "inner_steps_per_launch": args.inner,
                "parallelism": "env-index sharding x%d, no data-path collective" % world,
                "l2": "working set %.2f GB per GPU (> 126 MB L2); no flush needed" %
                      (n * (17216 + 128 + 192 + 18 + 4 * 288 + 2304) / 1e9)})
    return cfg


def peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def committed_profile():
    """profiles/traffic.json: per-kernel numbers taken from the committed ncu captures (DRAM bytes and warp
    instructions per env-step / per reset) — the part of the roofline that cannot be measured live."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        return {}


def issue_frac(units_per_s_per_gpu, warp_inst_per_unit, sm_mhz):
    """fraction of the GPU's warp-instruction issue slots (148 SMs x 4 schedulers x clock) the measured rate needs"""
    if not warp_inst_per_unit:
        return None
    return units_per_s_per_gpu * warp_inst_per_unit / (N_SMS * SMSP_PER_SM * sm_mhz * 1e6)


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region (profiling recipe's clocks line).  Read through
    NVML inside this process (one nvmlInit before the timed region, then two cheap queries every 20 ms): a
    fresh `nvidia-smi` process per sample attaches to the driver each time and stalls the GPU for 8-15 ms —
    measured here as 5.23e9 env-steps/s without sampling against 4.5-5.2e9 with it (profiles/r02_o_clock_sampling.txt).
    Falls back to nvidia-smi (one sample every 0.5 s) when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()
        self.t_begin = 0.0  # samples taken before mark_begin() (warm-up) are dropped
        self.nvml, self.handle, self.source = None, None, "nvidia-smi"
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            p = torch.cuda.get_device_properties(index)
            bus = "%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
            self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml, self.source = pynvml, "nvml"
            self._sample_nvml()  # the first query of a process is the slow one
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
        bits = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        flags = [bool(bits & n.nvmlClocksThrottleReasonHwSlowdown), bool(bits & n.nvmlClocksThrottleReasonHwThermalSlowdown),
                 bool(bits & n.nvmlClocksThrottleReasonSwThermalSlowdown), bool(bits & n.nvmlClocksThrottleReasonSwPowerCap)]
        return [mhz, self.max_mhz] + flags

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        r = [x.strip() for x in out.strip().split(",")]
        return [float(r[0]), float(r[1])] + [v.lower().startswith("active") for v in r[2:6]]

    def run(self):
        while not self._halt.is_set():
            try:
                t = time.perf_counter()
                self.rows.append((t, self._sample_nvml() if self.nvml else self._sample_smi()))
            except Exception:
                pass
            self._halt.wait(0.02 if self.nvml else 0.5)

    def mark_begin(self):
        """the timed region starts now (the thread has been polling since before the warm-up, so that its
        start-up and NVML's first queries fall outside the region)"""
        self.t_begin = time.perf_counter()

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for t, r in self.rows:
            if t < self.t_begin:
                continue
            sm.append(r[0]); mx = max(mx, r[1])
            reasons.update(name for name, on in zip(self.NAMES, r[2:6]) if on)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}


class Ctx:
    """rank bookkeeping: barrier + synchronize on both sides of a timed region, max over ranks"""

    def __init__(self, args, rank, local_rank, world):
        self.args, self.rank, self.local_rank, self.world = args, rank, local_rank, world

    def barrier(self):
        import torch
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, x):
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed_ms(self, fn, iters):
        """device time of `iters` calls of fn on the current stream, max over ranks"""
        import torch
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        a0.record()
        for _ in range(iters):
            fn()
        a1.record()
        self.barrier()
        return self.max_over_ranks(a0.elapsed_time(a1))


def roofline_fused(n, inner, k_ms, sm_mhz, stats):
    peak, peak_src = peak_hbm()
    prof = committed_profile().get("k_step_fused", {})
    algo_bytes = ALGO_BYTES_PER_ENV_STEP * n * inner
    achieved = algo_bytes / (k_ms * 1e-3) / 1e9
    traffic = None
    if prof.get("inner") == inner and prof.get("envs") == n:
        traffic = prof.get("dram_bytes_per_launch")
    rate = n * inner / (k_ms * 1e-3)
    winst = prof.get("warp_inst_per_env_step")
    return {"bound": "hbm", "kernel": "cog::k_step_fused<false,false> (+ k_step_tail for finished episodes)",
            "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": k_ms,
            "binding_resource": "warp-instruction issue / ALU pipe, not HBM",
            "issue_frac": issue_frac(rate, winst, sm_mhz), "warp_inst_per_env_step": winst,
            "alu_pipe_frac_ncu": prof.get("alu_pipe_frac"), "issue_active_frac_ncu": prof.get("issue_active_frac"),
            "dram_measured_GBs": (traffic / (k_ms * 1e-3) / 1e9) if traffic else None,
            "note": "frac is the NOMINAL fraction SURVEY 8(d) defines: 696 B per env-step (what would move if every "
                    "exported buffer were refreshed every step) x envs x inner / kernel time / measured HBM peak. The "
                    "fused kernel keeps env state in registers for the whole launch, so the bytes never move (traffic "
                    "= real DRAM bytes per launch from the committed ncu capture) and the kernel is bound by "
                    "instruction issue: issue_frac = warp instructions per env-step (ncu) x env-steps/s / (148 SMs x "
                    "4 schedulers x SM clock). The *_ncu fields and the per-step instruction count come from the "
                    "committed capture (profiles/r02_final_fused.txt), taken before the envs were handed out by "
                    "ticket: same instructions per env-step, but the 0.79 issue_active of that capture is the static "
                    "split's; issue_frac above is computed from this run's own rate"}


# --------------------------------------------------------------------------- device-side extras
def measure_resets(ctx, vec, n):
    """procedural board generation: full-batch resets (the reference's `time_reset` analogue)"""
    peak, _ = peak_hbm()
    for _ in range(2):
        vec.reset()
    r_iters = 10
    r_ms = ctx.timed_ms(vec.reset, r_iters) / r_iters
    per_gpu = n / (r_ms * 1e-3)
    prof = committed_profile().get("k_reset", {})
    gbs = ALGO_BYTES_PER_RESET * per_gpu / 1e9
    return {"value": ctx.world * per_gpu, "unit": "resets/s", "ms_per_batch_reset": r_ms, "kernel": "cog::k_reset",
            "algorithmic_bytes_per_reset": ALGO_BYTES_PER_RESET, "achieved_GBs_per_gpu": gbs, "frac": gbs / peak,
            "warp_inst_per_reset": prof.get("warp_inst_per_reset"),
            "dram_bytes_per_reset": prof.get("dram_bytes_per_reset"),
            "note": "vec.reset() of all %d envs per GPU, %d timed launches; frac = 19 440 B per reset (SURVEY 8d) "
                    "against the measured HBM peak" % (n, r_iters)}


def measure_side_workloads(ctx, L, n, offset, sm_mhz):
    """The other single-GPU workloads SURVEY §8(d) asks for (same kernels; the headline stays configs[2]
    regime A): regime M = actions sampled from the merged legal mask (moves, purchases, wins: every branch
    of step live), the reset-heavy configs[3], and the rollout writer."""
    from gym_eldorado_b200 import capi, layout, shard
    W = ctx.args.workload
    peak, _ = peak_hbm()
    prof = committed_profile()
    out = {}
    for name, diffc, npl, max_steps, merged, slots, pkey in (
            ("regime_M_merged_masks", W["difficulty"], 4, 100000, True, 0, "k_step_fused_M"),
            ("reset_heavy_configs3", "HARD", 2, 8, False, 0, None),
            ("reset_heavy_4_players", "HARD", 4, 8, False, 0, None),
            ("rollout_ring", W["difficulty"], 4, 100000, False, 256, None),
            ("rollout_ring_regime_M", W["difficulty"], 4, 100000, True, 256, None)):
        v2 = capi.VecHandle(n, offset, ctx.local_rank, None, L)
        s2 = capi.SamplerHandle(n, W["seed"], offset, ctx.local_rank, None, L)
        v2.reset(W["seed"], npl, 3, DIFFICULTY[diffc], max_steps)
        if slots:
            v2.rollout_attach(slots)
        inner2, reps = 256, 5
        for _ in range(3):
            v2.run_fused(s2, inner2, merged=merged)
        v2.stats_reset()
        sec = ctx.timed_ms(lambda: v2.run_fused(s2, inner2, merged=merged), reps) * 1e-3
        st2 = shard.reduce_stats(v2.stats(), device="cuda")
        rate = st2["env_steps"] / sec
        per_gpu = rate / ctx.world
        frac = ALGO_BYTES_PER_ENV_STEP * per_gpu / 1e9 / peak
        winst = (prof.get(pkey) or {}).get("warp_inst_per_env_step") if pkey else None
        out[name] = {"value": rate, "unit": "env-steps/s", "frac": frac,
                     "issue_frac": issue_frac(per_gpu, winst, sm_mhz), "warp_inst_per_env_step": winst,
                     "resets_per_s": st2["resets"] / sec, "episodes": st2["episodes"],
                     "wins": sum(st2["wins_seat%d" % k] for k in range(4)), "truncations": st2["truncations"],
                     "travelled_hexes": st2["travelled_hexes"], "cards_added": st2["cards_added"],
                     "config": "%d envs per GPU, %d players, %s, max_steps %d, %d launches x %d iterations, %s" % (
                         n, npl, diffc, max_steps, reps, inner2,
                         "merged legal masks" if merged else "selected_action_masks")}
        if name == "regime_M_merged_masks":
            out[name]["summary"] = {"value": rate, "unit": "env-steps/s", "roofline": {"bound": "hbm", "frac": frac,
                                    "issue_frac": out[name]["issue_frac"]}, "episodes": st2["episodes"],
                                    "travelled_hexes": st2["travelled_hexes"], "cards_added": st2["cards_added"]}
        if slots:
            out[name]["ring"] = "%d slots x %d envs x %d B per GPU" % (slots, n, layout.compact_dtype.itemsize)
            out[name]["ring_write_GBs_per_gpu"] = per_gpu * layout.compact_dtype.itemsize / 1e9
        v2.close(); s2.close()
    return out


def measure_api_loop(ctx, vec, smp, n):
    """The reference's own Python loop on DEVICE buffers (nothing crosses PCIe), per step: as two kernels
    (sampler.sample(envs.selected_action_masks); envs.step(sampler.get_actions())) and as one
    (cog_vec_step_sampled — what Runner.sample(); Runner.step() launch)."""
    from gym_eldorado_b200 import layout
    peak, _ = peak_hbm()
    masks_ptr = vec.buffer(layout.BUF_SELECTED_MASKS)[0]
    act_ptr = smp.actions_ptr()
    it = 200
    out = {}

    def two():
        smp.sample_device(masks_ptr); vec.step_device(act_ptr)

    def one():
        vec.step_sampled(smp)
    for name, fn, api in (("device_api_loop_two_kernels", two, "k_sample + k_step_single<false> per iteration"),
                          ("device_api_loop", one, "cog_vec_step_sampled: k_step_single<true>, one kernel per iteration")):
        for _ in range(10):
            fn()
        ms = ctx.timed_ms(fn, it)
        per_gpu = n * it / (ms * 1e-3)
        out[name] = {"value": ctx.world * per_gpu, "unit": "env-steps/s", "us_per_iteration": ms * 1e3 / it,
                     "frac": ALGO_BYTES_PER_ENV_STEP * per_gpu / 1e9 / peak, "api": api + ", %d iterations" % it}
    prof = committed_profile().get("k_step_single", {})
    out["device_api_loop"]["dram_bytes_per_env_step"] = prof.get("dram_bytes_per_env_step")
    out["device_api_loop"]["warp_inst_per_env_step"] = prof.get("warp_inst_per_env_step")
    return out


def measure_config4(ctx, L):
    """BASELINE configs[4] in the multi-GPU line: 1 048 576 envs split evenly over the ranks (strong scaling),
    4 players, HARD, fused kernel; the episode statistics are all-reduced over NCCL."""
    from gym_eldorado_b200 import capi, shard
    W4 = WORKLOADS[4]
    peak, _ = peak_hbm()
    lo, hi = shard.shard_range(ctx.rank, ctx.world, W4["n_envs_total"])
    n4 = hi - lo
    v = capi.VecHandle(n4, lo, ctx.local_rank, None, L)
    s = capi.SamplerHandle(n4, W4["seed"], lo, ctx.local_rank, None, L)
    v.reset(W4["seed"], W4["n_players"], W4["n_pieces"], DIFFICULTY[W4["difficulty"]], W4["max_steps"])
    inner, reps = 256, 4
    for _ in range(3):
        v.run_fused(s, inner)
    v.stats_reset()
    sec = ctx.timed_ms(lambda: v.run_fused(s, inner), reps) * 1e-3
    st = shard.reduce_stats(v.stats(), device="cuda")
    assert st["env_steps"] == W4["n_envs_total"] * inner * reps, "configs[4]: work was skipped"
    rate = st["env_steps"] / sec
    r_ms = ctx.timed_ms(v.reset, 3) / 3
    v.close(); s.close()
    return {"metric": METRIC, "value": rate, "unit": "env-steps/s", "scaling": "strong", "n_gpus": ctx.world,
            "frac_per_gpu": ALGO_BYTES_PER_ENV_STEP * rate / ctx.world / 1e9 / peak,
            "resets_per_s": W4["n_envs_total"] / (r_ms * 1e-3),
            "config": dict(W4, n_envs_per_gpu=n4, inner_steps_per_launch=inner, launches=reps),
            "episode_stats": st,
            "note": "1 M envs / %d ranks = %d per GPU; stats all-reduced over NCCL (96 bytes); no data-path collective"
                    % (ctx.world, n4)}


# --------------------------------------------------------------------------- end to end, host arrays
def _pinned(count_bytes):
    import torch
    return torch.zeros(count_bytes, dtype=torch.uint8).pin_memory()


def host_memcpy_gbs(nbytes=256 << 20):
    """host memory bandwidth as one thread sees it (numpy copy of a buffer far larger than the caches, read +
    write bytes) — what the staging copies and the zero-copy kernels of several ranks on one socket share"""
    import numpy as np
    a = np.ones(nbytes, dtype=np.uint8)
    b = np.empty_like(a)
    b[...] = a
    t0 = time.perf_counter()
    for _ in range(3):
        b[...] = a
    return 2 * 3 * nbytes / (time.perf_counter() - t0) / 1e9


def measure_e2e(ctx, L, vec, smp, n, steps):
    """Host policy loop through the C ABI: every iteration hands the library that step's inputs in HOST arrays
    (masks for sample(), actions for step()) and gets that step's results back in host arrays — pinned host
    memory, copies inside the timed region, wall clock with a barrier + synchronize on both sides.
      packed        cog_sampler_sample_packed_host_io + cog_vec_step_packed_host_io: 16-byte masks, 8-byte actions
                    (the headline `value`: the same information in 66 instead of 246-402 bytes per env-step)
      compact       the packed loop + the 256-byte compact observation of every env each step
      reference     the reference's own record layouts (ActionMask 128 B, ActionData 64 B): default path
                    (auto-selected), zero copy and DMA pipeline forced, and one blocking call per copy"""
    import numpy as np
    from gym_eldorado_b200 import layout
    mdt, adt = layout.action_mask_dtype, layout.action_dtype
    pm_dt, pa_dt = layout.packed_mask_dtype, layout.packed_action_dtype
    from gym_eldorado_b200 import shard
    with shard.near_gpu(ctx.local_rank) as numa:  # the rank's host arrays on the NUMA node of its GPU
        h_masks, h_actions = _pinned(n * mdt.itemsize), _pinned(n * adt.itemsize)
        h_pm, h_pa = _pinned(n * pm_dt.itemsize), _pinned(n * pa_dt.itemsize)
        h_rew, h_done, h_agent = _pinned(n * 16), _pinned(n), _pinned(n)
        h_compact = _pinned(n * layout.compact_dtype.itemsize)
    np_masks, np_actions = h_masks.numpy().view(mdt), h_actions.numpy().view(adt)
    np_pm, np_pa = h_pm.numpy().view(pm_dt), h_pa.numpy().view(pa_dt)
    np_rew = h_rew.numpy().view(np.float32).reshape(n, 4)
    np_done, np_agent = h_done.numpy().view(np.bool_), h_agent.numpy()
    np_compact = h_compact.numpy().view(layout.compact_dtype)

    def step_serial():  # one blocking call per copy / kernel (the round-1 path)
        smp.sample_host(np_masks)
        smp.read_actions(np_actions)
        vec.step_host(np_actions)
        vec.read(layout.BUF_SELECTED_MASKS, np_masks)
        L.cog_vec_read(vec._h, layout.BUF_REWARDS, h_rew.data_ptr(), h_rew.numel())
        L.cog_vec_read(vec._h, layout.BUF_DONES, h_done.data_ptr(), h_done.numel())
        L.cog_vec_read(vec._h, layout.BUF_AGENT_SELECTION, h_agent.data_ptr(), h_agent.numel())

    def step_reference_layout():
        smp.sample_host_io(np_masks, np_actions)
        vec.step_host_io(np_actions, np_masks, np_rew, np_done, np_agent)

    def step_packed():
        smp.sample_packed_host_io(np_pm, np_pa)
        vec.step_packed_host_io(np_pa, np_pm, np_rew, np_done, np_agent)

    def step_compact():
        step_packed()
        vec.compact_obs_device(None, merged=False)
        vec.read(layout.BUF_COMPACT, np_compact)

    def timed(fn, iters, warm=12):
        for _ in range(warm):  # also covers the library's nine-call measurement of its two host-I/O paths
            fn()
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        ctx.barrier()
        return ctx.world * n * iters / ctx.max_over_ranks(time.perf_counter() - t0)

    iters = max(30, steps * 5)
    vec.packed_masks(out=np_pm)
    packed = timed(step_packed, iters)
    compact = timed(step_compact, max(10, iters // 3), warm=3)
    vec.read(layout.BUF_SELECTED_MASKS, np_masks)
    serial = timed(step_serial, max(10, iters // 3), warm=3)
    vec.read(layout.BUF_SELECTED_MASKS, np_masks)
    os.environ["COG_IO_ZEROCOPY"] = "0"
    piped = timed(step_reference_layout, iters, warm=3)
    os.environ["COG_IO_ZEROCOPY"] = "1"
    vec.read(layout.BUF_SELECTED_MASKS, np_masks)
    zero = timed(step_reference_layout, iters, warm=3)
    os.environ.pop("COG_IO_ZEROCOPY")
    vec.read(layout.BUF_SELECTED_MASKS, np_masks)
    direct = timed(step_reference_layout, iters)  # the default: the library measures both paths, keeps the faster
    io_path = vec.io_path(smp)
    h2d = n * (pm_dt.itemsize + pa_dt.itemsize)
    d2h = n * (pa_dt.itemsize + pm_dt.itemsize + 16 + 1 + 1)
    full = n * (mdt.itemsize + adt.itemsize)
    return {"value": packed, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "api": "host policy loop, every step: cog_sampler_sample_packed_host_io(host packed masks -> host packed "
                   "actions); cog_vec_step_packed_host_io(host packed actions -> host packed masks, rewards, dones, "
                   "agent_selection); single-step kernels; pinned host arrays staged by DMA; %d iterations" % iters,
            "compact_value": compact,
            "compact_d2h_bytes_per_step": d2h + n * layout.compact_dtype.itemsize,
            "reference_layout_value": direct, "reference_layout_io_path": {"env": io_path[0], "sampler": io_path[1]},
            "reference_layout_bytes_per_step": {"zero_copy": {"h2d": n * (96 + 32), "d2h": n * (92 + 8 + 18)},
                                                "dma": {"h2d": full, "d2h": full + n * 18}},
            "zero_copy_value": zero, "dma_pipeline_value": piped, "serial_calls_value": serial,
            "host_memcpy_GBs": host_memcpy_gbs() if ctx.rank == 0 else None,
            "host_arrays_numa": numa.info,
            "note": "value = packed records (opt-in entry points: the same information as the reference's ActionMask / "
                    "ActionData in 16 + 8 bytes); reference_layout_value = the reference's own record layouts through "
                    "cog_sampler_sample_host_io / cog_vec_step_host_io (path auto-selected per handle: "
                    "reference_layout_io_path), zero_copy_value / dma_pipeline_value = that loop with COG_IO_ZEROCOPY=1 / 0, "
                    "serial_calls_value = one blocking call per copy / kernel; compact_value = the packed loop plus the "
                    "256-byte compact observation of every env copied to the host each step. All GPUs of this box sit on "
                    "one NUMA node (nvidia-smi topo: CPU affinity 0-31 for every GPU), so there is no per-rank placement "
                    "to choose; host_memcpy_GBs is what the ranks share"}


# --------------------------------------------------------------------------- the reference on the host cores
def _host_cores():
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    return cores


def _ref_setup(args, n_envs=None, n_players=None, difficulty=None, max_steps=None):
    """(vec, kind, n): the unmodified reference (oracle/_ref) where it was built, else the C port"""
    from oracle import refshim
    W = args.workload
    n = n_envs or args.envs
    diff = DIFFICULTY[difficulty or W["difficulty"]]
    npl = n_players or W["n_players"]
    ms = W["max_steps"] if max_steps is None else max_steps
    if refshim.available():
        n = max(k for k in refshim.SUPPORTED_N if k <= n)  # the reference's templates are instantiated for these sizes
        if npl < 4:
            npl = 4  # fewer than 4 players is undefined behaviour in the reference
        v = refshim.RefVec(n, W["seed"])
        v.reset(W["seed"], npl, 3, diff, ms)
        return v, "reference", n
    from oracle import port
    v = port.OracleVec(n, W["seed"])
    v.reset(W["seed"], npl, 3, diff, ms)
    return v, "port", n


def _ref_runner(args):
    """the headline CPU arm: ThreadedRunner<N> async with every usable host thread but one"""
    v, kind, n = _ref_setup(args)
    cores = _host_cores()
    if kind == "reference":
        threads = max(1, min(cores - 1, 32))  # workers busy-wait and pin; never oversubscribe
        v.runner(threads)
        return v, kind, n, threads, (lambda k: v.time_run(k, 1)), \
            "unmodified reference ThreadedRunner<%d> async, %d pinned worker threads" % (n, threads)
    return v, kind, n, 1, v.time_run, "oracle port (scalar C), 1 thread"


def cpu_baseline(args, bounded_seconds=10.0):
    v, kind, n, used, run, how = _ref_runner(args)
    t = run(2)
    inner = max(1, int(2 * bounded_seconds / max(t, 1e-6)))
    dt = run(inner)
    v.close()
    return {"value": n * inner / dt, "unit": "env-steps/s", "cores": used, "kind": kind,
            "sample": "%s; %d sample+step iterations on %d envs (%.1f s)" % (how, inner, n, dt)}


def cpu_legs(args, seconds=1.2):
    """The CPU baselines BASELINE.md §3 lists (reference benchmarks/benchmarks.py:47-63), each a bounded sample
    on this box: thread sweep async + sync at N = 256 and 65 536, the sampler alone, sequential resets, and the
    reset-heavy workload sequentially (threaded resets race on the reference's global piece tables)."""
    from oracle import refshim
    if not refshim.available():
        return {"unavailable": "oracle/_ref not built"}
    cores = _host_cores()
    W = args.workload

    def rate(run, n):
        t = run(2)
        k = max(2, int(2 * seconds / max(t, 1e-6)))
        return n * k / run(k)
    out = {"cores": cores, "unit": "env-steps/s", "sweep": {}}
    for n in (256, 65536):
        v = refshim.RefVec(n, W["seed"])
        v.reset(W["seed"], 4, 3, DIFFICULTY[W["difficulty"]], W["max_steps"])
        # every measurement starts from a fresh reset: at N = 256 a few seconds of stepping reach max_steps, and the
        # auto-resets that follow race on the reference's global piece tables under the threaded runner
        fresh = lambda: v.reset(W["seed"], 4, 3, DIFFICULTY[W["difficulty"]], W["max_steps"])
        row = {"sequential": rate(lambda k: v.time_run(k, 0), n)}
        for T in sorted({t for t in (1, 2, 4, 8, 16, cores - 1) if 1 <= t <= max(1, cores - 1)}):
            row["T%d" % T] = {}
            for mode, name in ((1, "async"), (2, "sync")):
                fresh()
                v.runner(T)
                row["T%d" % T][name] = rate(lambda k: v.time_run(k, mode), n)
                v.drop_runner()
        fresh()
        if n == 65536:
            t = v.time_sample(2)
            k = max(2, int(2 * seconds / max(t, 1e-6)))
            out["sample_only"] = {"value": n * k / v.time_sample(k), "unit": "env-samples/s", "threads": 1}
        out["sweep"]["N%d" % n] = row
        v.close()
    # resets: 4 096 envs, sequential (benchmarks.py time_reset)
    v = refshim.RefVec(4096, W["seed"])
    for name, diffc in (("MEDIUM", 1), ("HARD", 2)):
        v.reset(W["seed"], 4, 3, diffc, 100000)
        dt = v.time_reset(1)
        out.setdefault("resets", {})[name] = {"value": 4096 / dt, "unit": "resets/s", "threads": 1}
    v.close()
    # reset-heavy workload (configs[3] twin at 4 players), sequential
    v = refshim.RefVec(4096, W["seed"])
    v.reset(W["seed"], 4, 3, DIFFICULTY["HARD"], 8)
    out["reset_heavy_4_players_sequential"] = {"value": rate(lambda k: v.time_run(k, 0), 4096), "unit": "env-steps/s",
                                               "threads": 1, "config": "4 096 envs, 4 players, HARD, max_steps 8"}
    v.close()
    return out


def reference_arm(args, world):
    """bench.py --impl reference: the reference's own CPU implementation of the path on this box's host cores,
    same config / metric / unit; each step a bounded sample of the workload."""
    v, kind, n, used, run, how = _ref_runner(args)
    t = run(2)
    # about one second per step, less when many steps are asked for: the whole run stays near 30 s
    per_step_s = max(0.2, min(1.0, 30.0 / max(1, args.steps + args.warmup)))
    ref_inner = max(1, int(2 * per_step_s / max(t, 1e-6)))
    for _ in range(args.warmup):
        run(ref_inner)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(ref_inner)
    dt = time.perf_counter() - t0
    v.close()
    value = n * ref_inner * args.steps / dt
    sample = "%s; each step = %d sample+step iterations on %d envs (bounded sample of the workload)" % (how, ref_inner, n)
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "strong" if args.workload.get("n_envs_total") else "weak",
            "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": config_of(args, world),
            "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": used, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
