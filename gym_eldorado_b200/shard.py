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


# ---- host memory next to the GPU -------------------------------------------------------------------
# On a multi-socket box the host arrays of a rank's policy loop should live on the NUMA node its GPU hangs off:
# pinned memory is placed where the allocating thread runs (first touch), and with eight ranks on one host
# the far socket's memory costs every PCIe transfer a hop over the socket interconnect.
def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def numa_cpus_of_pci(bus_id, sysfs="/sys"):
    """(node, set of CPU ids) of the NUMA node a PCI device ("0000:1b:00.0") is attached to, or None when the
    platform does not say (single node, container without sysfs, unknown device)."""
    import os
    try:
        node = int(open(os.path.join(sysfs, "bus/pci/devices", bus_id.lower(), "numa_node")).read())
        if node < 0:
            return None
        cpus = _parse_cpulist(open(os.path.join(sysfs, "devices/system/node/node%d/cpulist" % node)).read())
        return (node, cpus) if cpus else None
    except (OSError, ValueError):
        return None


class near_gpu:
    """Context manager: run the calling thread on the CPUs of the NUMA node of CUDA device `device` (allocate
    the rank's pinned host arrays inside it), restore the previous affinity on exit.  `info` says what was
    done; a no-op wherever the platform gives no answer."""

    def __init__(self, device):
        self.device, self.info, self._old = device, {"numa_node": None, "bound": False}, None

    def __enter__(self):
        import os
        try:
            p = torch.cuda.get_device_properties(self.device)
            bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
            self.info["pci"] = bus
            found = numa_cpus_of_pci(bus)
            if found:
                node, cpus = found
                self._old = os.sched_getaffinity(0)
                usable = cpus & self._old
                if usable:
                    os.sched_setaffinity(0, usable)
                    self.info.update({"numa_node": node, "bound": True, "cpus": len(usable)})
        except Exception as e:  # placement is an optimisation, never a failure
            self.info["error"] = repr(e)
        return self

    def __exit__(self, *exc):
        import os
        if self._old and self.info.get("bound"):
            os.sched_setaffinity(0, self._old)
        return False
