"""Host-side partitioning of the IK path over ranks (one process per GPU).

Grid points are independent given the whole frame, and frames are independent of each other -- exactly the
reference's MPI decomposition (pyglasstools/irvingkirkwood/__init__.py:53-60: ``np.array_split`` chunks of the
i-major grid, one per rank; every rank reads the full frame).  Nothing here touches a GPU: the same functions
drive the NCCL ranks of bench.py and the gloo ranks of the CPU tests.  (The torch.distributed helpers of the test harness
live in tests/dist_helpers.py: the product package does not import torch.)
"""
import numpy as np


def grid_shards(total, world):
    """[g0, g1) per rank, identical to ``np.array_split(np.arange(total), world)`` boundaries
    (the first ``total % world`` shards are one longer)."""
    base, extra = divmod(int(total), int(world))
    sizes = np.full(world, base, dtype=np.int64)
    sizes[:extra] += 1
    ends = np.cumsum(sizes)
    return [(int(e - s), int(e)) for s, e in zip(sizes, ends)]


def frame_shards(nframes, world, rank):
    """Frames of a trajectory owned by ``rank`` when frames are dealt round-robin."""
    return list(range(int(rank), int(nframes), int(world)))
