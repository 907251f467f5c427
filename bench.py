#!/usr/bin/env python
"""bench.py — env-steps/s of the City of Gold vectorised hot path on B200.

One bench "step" = one launch of the persistent fused kernel: `--inner` iterations of
{masked random sample; game step (+auto-reset)} on every env of the workload
(BASELINE.json configs[2]: 65 536 envs per GPU, 4 players, MEDIUM).  Prints ONE JSON
line (see the contract in the task description / DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

The line's parts (bench_parts.py holds the side measurements):
  value / roofline     the fused kernel, device resident (regime A = the reference's own benchmark flow)
  extra                regime M (every branch of step live), reset-heavy configs[3], resets/s, the
                       per-step device API loop (two kernels and one), rollout ring — each with its own
                       fraction of the SURVEY 8(d) roofline
  config4              (N > 1) BASELINE configs[4]: 1 M envs split over the ranks, HARD, strong scaling
  e2e                  host policy loop through the C ABI with host arrays, copies inside the timed region
  cpu_baseline, cpu_legs   the unmodified reference on this box's host cores
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import bench_parts as bp  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--inner", type=int, default=1024, help="fused sample+step iterations per launch")
    ap.add_argument("--config", type=int, default=2, choices=sorted(bp.WORKLOADS), help="BASELINE.json configs index")
    ap.add_argument("--envs", type=int, default=None, help="envs per GPU (default: the config's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cpu-legs", action="store_true", help="skip the thread sweep / reset / sample CPU legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-clocks", action="store_true", help="do not poll nvidia-smi during the timed region (diagnosis)")
    ap.add_argument("--no-resets", action="store_true")
    ap.add_argument("--no-side", action="store_true", help="skip the regime-M / reset-heavy / API-loop side measurements")
    ap.add_argument("--no-config4", action="store_true", help="skip the configs[4] leg of a multi-GPU run")
    args = ap.parse_args()
    args.workload = dict(bp.WORKLOADS[args.config])
    if args.envs is None:
        args.envs = args.workload["n_envs_per_gpu"] or args.workload["n_envs_total"] // max(1, args.gpus)
    args.workload["n_envs_per_gpu"] = args.envs
    return args


def main():
    args = parse()
    # ONE line on stdout: whatever libraries print there while the job runs (NCCL's version banner under
    # torchrun, for one) goes to stderr; the JSON line is written to the real stdout at the end
    sys.stdout.flush()
    out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(bp.reference_arm(args, world)), file=out, flush=True)
        return

    import torch
    import torch.distributed as dist
    from gym_eldorado_b200 import capi, shard

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = bp.Ctx(args, rank, local_rank, world)
    W = args.workload
    L = capi.lib()
    n = args.envs
    offset = rank * n  # weak scaling: every GPU owns its own slice of global env indices
    if W.get("n_envs_total") and args.envs * world == W["n_envs_total"]:
        offset, stop = shard.shard_range(rank, world, W["n_envs_total"])  # configs[4]: fixed total
        n = stop - offset
    vec = capi.VecHandle(n, offset, local_rank, None, L)
    smp = capi.SamplerHandle(n, W["seed"], offset, local_rank, None, L)
    vec.reset(W["seed"], W["n_players"], W["n_pieces"], bp.DIFFICULTY[W["difficulty"]], W["max_steps"])
    launches0 = L.cog_kernel_launches()

    # ---- the timed region: K launches of the fused kernel --------------------------------------
    clocks = bp.ClockSampler(local_rank)  # attaches to NVML and starts polling now, well before the timed region
    if rank == 0 and not args.no_clocks:
        clocks.start()
    for _ in range(args.warmup):
        vec.run_fused(smp, args.inner)
    ctx.barrier()
    clocks.mark_begin()
    vec.stats_reset()
    launches_timed0 = L.cog_kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        vec.run_fused(smp, args.inner)
    # the one data-path-free collective: aggregate episode statistics over NVLink
    stats = shard.reduce_stats(vec.stats(), device="cuda")
    ev1.record()
    ctx.barrier()
    ms = ctx.max_over_ranks(ev0.elapsed_time(ev1))
    clk = clocks.finish() if rank == 0 and not args.no_clocks else None
    launches_timed = L.cog_kernel_launches() - launches_timed0
    assert stats["env_steps"] == world * n * args.inner * args.steps, "work was skipped"
    value = stats["env_steps"] / (ms * 1e-3)

    # per-launch duration of the dominant kernel, CUDA events recorded by the library on its own stream
    per_launch = []
    for _ in range(max(3, min(args.steps, 10))):
        vec.run_fused(smp, args.inner)
        per_launch.append(vec.last_kernel_ms())
    k_ms = sum(per_launch) / len(per_launch)
    sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
    roofline = bp.roofline_fused(n, args.inner, k_ms, sm_mhz, stats)

    extra = {}
    if not args.no_resets:
        extra["resets"] = bp.measure_resets(ctx, vec, n)
    if not args.no_side:
        extra.update(bp.measure_side_workloads(ctx, L, n, offset, sm_mhz))
        extra.update(bp.measure_api_loop(ctx, vec, smp, n))
    config4 = None
    if world > 1 and not args.no_config4 and args.config == 2:
        config4 = bp.measure_config4(ctx, L)
    e2e = None if args.no_e2e else bp.measure_e2e(ctx, L, vec, smp, n, args.steps)
    launches = L.cog_kernel_launches() - launches0

    cpu, legs = None, None
    if rank == 0 and not args.no_cpu_baseline:
        try:
            cpu = bp.cpu_baseline(args, bounded_seconds=10.0)
            if not args.no_cpu_legs and world == 1:
                legs = bp.cpu_legs(args)
        except Exception as e:  # the baseline is reporting, never the product path
            cpu = cpu or {"error": repr(e)}
            legs = {"error": repr(e)}

    if rank == 0:
        line = {"metric": bp.METRIC, "value": value, "unit": "env-steps/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "strong" if W.get("n_envs_total") else "weak",
                "vs_baseline": None, "dtype": bp.DTYPE, "data": "synthetic",
                "config": bp.config_of(args, world), "roofline": roofline,
                "regime_M": (extra.get("regime_M_merged_masks") or {}).get("summary"),
                "extra": extra, "config4": config4, "cpu_baseline": cpu, "cpu_legs": legs, "e2e": e2e,
                "gpu_launches": launches_timed, "gpu_launches_total": launches, "clocks": clk,
                "episode_stats": stats,
                "note": "headline = regime A, the reference's own README / pytest / asv flow (actions sampled from "
                        "selected_action_masks: only play / null ever occur, %d episodes finished in the timed region); "
                        "regime_M = the same kernels with every branch of step live" % stats["episodes"]}
        print(json.dumps(line), file=out, flush=True)
    vec.close()
    smp.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
