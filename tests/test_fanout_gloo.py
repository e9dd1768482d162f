"""CPU tests of the multi-GPU plumbing: shard arithmetic and the cross-rank folding of per-rank
timings, run as two real processes over the gloo backend (the N>1 path of bench.py without GPUs)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import __graft_entry__ as entry


def test_shard_ranges_partition_the_array(pkg):
    f = pkg.fanout
    for n in (0, 1, 7, 50000, (1 << 28) + 5):
        for world in (1, 2, 3, 4, 8):
            ranges = [f.shard_range(n, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        f.shard_range(10, 2, 2)


def test_rank_seeds_and_status_names(pkg):
    f = pkg.fanout
    seeds = [f.rank_seed(0xB200, r) for r in range(8)]
    assert len(set(seeds)) == 8 and seeds[0] == 0xB200
    assert f.status_file_names(2) == ["cuda-gpu0-ready", "cuda-gpu1-ready", "cuda-ready"]
    assert f.aggregate_gbs([3e9, 3e9], 0.5) == 12.0
    with pytest.raises(ValueError):
        f.aggregate_gbs([1.0], 0.0)
    # without an initialised process group the folds are the identity
    assert f.max_over_ranks(1.5) == 1.5 and f.sum_over_ranks(2.0) == 2.0


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    f = entry.load_package().fanout
    seconds = 0.25 * (rank + 1)                 # rank 1 is the slow one
    nbytes = 3.0e9
    dist.barrier()
    res = (f.max_over_ranks(seconds), f.sum_over_ranks(nbytes), f.job_throughput_gbs(nbytes, seconds),
           f.shard_range(101, rank, world), f.rank_seed(7, rank))
    torch.save(res, os.path.join(out, "r%d.pt" % rank))
    dist.destroy_process_group()


def test_cross_rank_folding_over_gloo(tmp_path):
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    assert r0[:3] == r1[:3]                      # every rank sees the same folded numbers
    slowest, total, gbs = r0[:3]
    assert slowest == 0.5 and total == 6.0e9 and gbs == 12.0
    assert r0[3] == (0, 51) and r1[3] == (51, 101)
    assert r0[4] != r1[4]
