"""torch.distributed helpers of the test harness (gloo on CPU): scalar reductions, whole-job throughput and a ragged
field gather.  Harness plumbing only -- the product's collectives are the NCCL entry points of include/pgt_cuda.h."""
import numpy as np

from pyglasstools_b200.sharding import grid_shards


def reduce_scalar(dist, value, op, device="cpu"):
    """max / sum of a Python float over the ranks of an initialised torch.distributed group."""
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op.upper()))
    return float(t.item())


def job_throughput(dist, units_this_rank, ms_this_rank, device="cpu"):
    """Whole-job units/s: all units of all ranks over the slowest rank's device time."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_this_rank / (ms_this_rank * 1e-3)
    total = reduce_scalar(dist, units_this_rank, "sum", device)
    worst = reduce_scalar(dist, ms_this_rank, "max", device)
    return total / (worst * 1e-3)


def gather_field(dist, local_field, total, device="cpu"):
    """Final field gather with torch.distributed (gloo on CPU tests; bench.py uses pgt_comm_allgather_field over
    NCCL): ragged shards are padded to the longest one, gathered, and trimmed in rank order."""
    import torch
    world = dist.get_world_size()
    shards = grid_shards(total, world)
    longest = max(e - s for s, e in shards)
    ncomp = local_field.shape[1]
    buf = torch.zeros((longest, ncomp), dtype=torch.float64, device=device)
    buf[: local_field.shape[0]] = torch.as_tensor(local_field, dtype=torch.float64, device=device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    return np.concatenate([p[: e - s].cpu().numpy() for p, (s, e) in zip(parts, shards)], axis=0)
