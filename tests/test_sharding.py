"""CPU, world_size 2 over gloo: the N>1 host logic — global-index sharding and the one
collective of the path (all-reduce of the episode statistics).  The kernels run under the
lock-step emulator (tests/emu); on the GPU box the same code runs on NCCL (bench.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity
from gym_eldorado_b200 import layout, shard

N_TOTAL, SEED, STEPS = 14, 4242, 90


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 16, 1 << 20):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(r, world, n) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(2, 2, 10)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lib = parity.emu_lib()
        lo, hi = shard.shard_range(rank, world, N_TOTAL)
        c = parity.Candidate(lib, hi - lo, SEED, offset=lo)
        c.reset(SEED, 4, 3, layout.HARD, 6)
        c.vec.run_fused(c.smp, STEPS)
        total = shard.reduce_stats(c.vec.stats())
        snap = c.snapshot()
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), lo=lo, hi=hi,
                 total=np.array([total[k] for k in sorted(total)], dtype=np.int64),
                 **{k: np.frombuffer(v.tobytes(), dtype=np.uint8) for k, v in snap.items()})
        c.close()
    finally:
        dist.destroy_process_group()


def test_two_ranks_equal_one_process(tmp_path, oracle_cls):
    parity.emu_lib()  # build once, before the workers race for it
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    # the same N_TOTAL envs in one process
    one = parity.Candidate(parity.emu_lib(), N_TOTAL, SEED)
    one.reset(SEED, 4, 3, layout.HARD, 6)
    one.vec.run_fused(one.smp, STEPS)
    ref_stats, ref = one.vec.stats(), one.snapshot()
    assert ref_stats["episodes"] > 0 and ref_stats["env_steps"] == N_TOTAL * STEPS
    # ... and the oracle
    o = oracle_cls(N_TOTAL, SEED)
    o.reset(SEED, 4, 3, layout.HARD, 6)
    for _ in range(STEPS):
        o.sample()
        o.step()
    assert parity.first_mismatch(ref, parity.oracle_snapshot(o)) is None
    for rank in range(world):
        z = np.load(os.path.join(str(tmp_path), "rank%d.npz" % rank))
        lo, hi = int(z["lo"]), int(z["hi"])
        assert list(z["total"]) == [ref_stats[k] for k in sorted(ref_stats)], "all-reduced statistics"
        for k, v in ref.items():  # every exported buffer of the shard == that slice of the single run
            assert z[k].tobytes() == v[lo:hi].tobytes(), (rank, k)
    one.close(); o.close()


def test_numa_lookup_of_a_pci_device(tmp_path):
    """shard.numa_cpus_of_pci reads the node and its CPU list from sysfs; unknown devices and single-node
    platforms (numa_node = -1) give None"""
    from gym_eldorado_b200 import shard
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("16-31,48-50\n")
    assert shard.numa_cpus_of_pci("0000:1B:00.0", str(tmp_path)) == (1, set(range(16, 32)) | {48, 49, 50})
    (dev / "numa_node").write_text("-1\n")
    assert shard.numa_cpus_of_pci("0000:1b:00.0", str(tmp_path)) is None
    assert shard.numa_cpus_of_pci("0000:ff:00.0", str(tmp_path)) is None
