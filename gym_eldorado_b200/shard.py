"""Multi-GPU plumbing of the path: environments are independent, so rank g of G
owns one contiguous slice of GLOBAL env indices and seeds derive from the global index —
any G replays identical per-env trajectories.  Nothing crosses GPUs on the data path; the
only collective is one all-reduce of the small episode-statistics vector (NCCL over NVLink on
the GPU box, gloo in the CPU tests)."""
import torch
import torch.distributed as dist

from . import capi


def shard_range(rank, world, n_total):
    """[start, stop) of the global env indices rank `rank` owns; slices differ by at most one env."""
    if not (0 <= rank < world) or n_total < 0:
        raise ValueError("bad rank/world/n_total")
    base, extra = divmod(n_total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def reduce_stats(local_stats, device=None, group=None):
    """Sum the per-rank episode statistics (dict from VecHandle.stats()) over all ranks.
    Returns the same dict of Python ints; a no-op without an initialised process group."""
    t = torch.tensor([int(local_stats[k]) for k in capi.STAT_NAMES], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, group=group)
    return dict(zip(capi.STAT_NAMES, (int(x) for x in t.tolist())))
