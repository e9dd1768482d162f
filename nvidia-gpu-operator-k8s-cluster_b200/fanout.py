"""Per-GPU fan-out arithmetic: the validation workload shards with no exchange.

Every rank (one process per GPU) validates its own device on its own arrays; nothing crosses
ranks on the data path.  These helpers only give every rank its own input seed and fold per-rank
results into the whole-job figure — bench.py uses every one of them; host/validator_core.cpp
implements the same folding for the `nvidia-validator` fan-out.
"""


def rank_seed(base_seed, rank):
    """Input seed of one rank's private arrays (weak scaling: every GPU gets its own data)."""
    return (base_seed + 0x1000 * rank) & (2**64 - 1)


def aggregate_gbs(bytes_per_rank, max_seconds):
    """Whole-job GB/s: bytes all ranks moved over the slowest rank's time."""
    if max_seconds <= 0:
        raise ValueError("non-positive time")
    return sum(bytes_per_rank) / max_seconds / 1e9


# ---- cross-rank folding (torch.distributed is plumbing only: a barrier and two reductions) -------

def _dist():
    import torch.distributed as dist
    return dist if (dist.is_available() and dist.is_initialized()) else None


def max_over_ranks(value, device="cpu"):
    """Max of a per-rank scalar over all ranks (the slowest rank's time).  Identity when
    torch.distributed is not initialised."""
    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device="cpu"):
    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def job_throughput_gbs(bytes_this_rank, seconds_this_rank, device="cpu"):
    """Whole-job GB/s from per-rank figures: all ranks' bytes over the slowest rank's time."""
    total = sum_over_ranks(bytes_this_rank, device)
    slowest = max_over_ranks(seconds_this_rank, device)
    return aggregate_gbs([total], slowest)
