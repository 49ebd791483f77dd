"""world_size-2 gloo test (CPU) of the multi-rank host logic: shard boundaries = the reference's np.array_split,
frames dealt round-robin, the max-over-ranks / sum-over-ranks throughput rule of bench.py, and the ragged field
gather.  The per-rank compute is done by the CPU oracle standing in for the GPU (same call shape: a shard
[g0,g1) of the i-major grid given the whole frame), so the assembled field must equal the single-rank one bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_grid_shards_equal_array_split():
    from pyglasstools_b200 import sharding
    for total in (0, 1, 5, 16129, 4194304):
        for world in (1, 2, 3, 4, 8):
            want = [len(c) for c in np.array_split(np.arange(total), world)]
            got = sharding.grid_shards(total, world)
            assert [e - s for s, e in got] == want
            assert got[0][0] == 0 and got[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
    assert sharding.frame_shards(10, 4, 1) == [1, 5, 9]
    assert sorted(sum((sharding.frame_shards(100, 8, r) for r in range(8)), [])) == list(range(100))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, outdir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from oracle import orc
    from pyglasstools_b200 import sharding
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import dist_helpers
    dist.init_process_group("gloo", rank=rank, world_size=world)
    z = np.load(os.path.join(ROOT, "tests", "golden", "polydisperse12_frames.npz"))
    pos = np.zeros((1024, 3), np.float32)
    pos[:, :2] = z["xy"][180]
    S = orc.System(z["box"], pos, z["diameter"])
    P, cg = orc.potential("poly12"), orc.cgfunc("octic", 2.0, 5)
    grid, _ = orc.cartesian_grid(float(z["box"][0]), 1.0)            # 31 x 31 = 961 points: ragged over 2 ranks
    g0, g1 = sharding.grid_shards(len(grid), world)[rank]
    local = S.ik(P, cg, grid[g0:g1])["virial"]
    full = dist_helpers.gather_field(dist, local, len(grid))
    # throughput rule: all units over the slowest rank
    value = dist_helpers.job_throughput(dist, g1 - g0, 10.0 * (rank + 1))
    frames = sharding.frame_shards(7, world, rank)
    nfr = dist_helpers.reduce_scalar(dist, len(frames), "sum")
    if rank == 0:
        np.savez(os.path.join(outdir, "out.npz"), full=full, value=value, nfr=nfr)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_grid_sharding_and_gather(tmp_path, orc, fixture_frames):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    out = np.load(str(tmp_path / "out.npz"))
    S = orc.System(fixture_frames.box, fixture_frames.position(180), fixture_frames.diameter)
    grid, _ = orc.cartesian_grid(float(fixture_frames.box[0]), 1.0)
    want = S.ik(orc.potential("poly12"), orc.cgfunc("octic", 2.0, 5), grid)["virial"]
    assert len(grid) == 961
    np.testing.assert_array_equal(out["full"], want)                 # bit-identical for every sharding
    assert float(out["value"]) == pytest.approx(961 / 20e-3)         # sum of units / max of times
    assert float(out["nfr"]) == 7
