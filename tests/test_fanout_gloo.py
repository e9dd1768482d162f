"""CPU tests of the multi-GPU plumbing: the per-rank seeds and the cross-rank folding of per-rank
timings that bench.py uses for `value` and `e2e`, run as two real processes over the gloo backend
(the N>1 path of bench.py without GPUs)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import __graft_entry__ as entry


def test_rank_seeds_and_aggregate(pkg):
    f = pkg.fanout
    seeds = [f.rank_seed(0xB200, r) for r in range(8)]
    assert len(set(seeds)) == 8 and seeds[0] == 0xB200
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
           f.aggregate_gbs([nbytes] * world, f.max_over_ranks(seconds)), f.rank_seed(7, rank))
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
    assert r0[3] == r1[3] == 12.0               # bench.py's `value` arithmetic: same on every rank
    assert r0[4] != r1[4]
