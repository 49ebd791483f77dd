"""Grid-point sharding over GPUs (the reference's only decomposition: np.array_split of the i-major grid over MPI ranks,
irvingkirkwood/__init__.py:45-63) through the C-ABI: slab-local stages 1-3, frame parts assembled with ncclAllGather, the
field gathered with NCCL, particle-sharded thermo sums reduced with ncclAllReduce.

One rank per GPU; inside one test process the ranks are threads (ctypes releases the GIL in the C-ABI calls).  With one
GPU the same entry points run as a one-rank communicator, so the NCCL code path is exercised on every GPU box.
"""
import threading

import numpy as np
import pytest
import torch  # noqa: F401  (before the library loads NCCL: torch bundles its own libnccl.so.2, see pgt_capi.cu load_nccl)

from pyglasstools_b200 import _capi, sharding, synth

pytestmark = pytest.mark.gpu

N_SIDE = 192          # 36 864 particles at rho = 1, box 192 x 192
NMAX = 192


def _frame():
    return synth.synth_frame(N_SIDE, 4242)


def _setup(ctx, fr, rc=1.25):
    ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    ctx.set_cg("octic", 2.0, 5)
    dmax = float(fr["diameter"].max())
    ctx.set_max_rcut(max(rc * dmax, dmax))
    axis = synth.linspace_axis(float(fr["box"][0]), NMAX)
    ctx.set_grid_cartesian(axis, axis)


def _full_field(fr, variant=0):
    ctx = _capi.Context(0)
    _setup(ctx, fr)
    ctx.set_option("ik_variant", variant)
    ctx.upload(fr["position"], fr["diameter"], fr["box"])
    ctx.build_cells()
    ctx.build_bonds()
    ctx.ik_compute(_capi.OBS_VIRIAL)
    full = ctx.ik_download()
    thermo = ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY)
    n_live = ctx.counts()["n_live"]
    ctx.close()
    return full, thermo, n_live


def test_slab_local_stages_reproduce_the_field_bit_for_bit():
    """A grid shard builds its cell and bond lists from the particles it can reach only; the shard's field is the
    corresponding slice of the whole-grid field, bit for bit, for both gather kernels."""
    fr = _frame()
    G = NMAX * NMAX
    ctx = _capi.Context(0)
    _setup(ctx, fr)
    for variant in (2, 0):
        full, _, n_all = _full_field(fr, variant)
        ctx.set_option("ik_variant", variant)
        for world in (2, 3, 8):
            kept = []
            for g0, g1 in sharding.grid_shards(G, world):
                ctx.set_shard(g0, g1)
                ctx.upload(fr["position"], fr["diameter"], fr["box"])
                ctx.build_cells()
                ctx.build_bonds()
                ctx.ik_compute(_capi.OBS_VIRIAL)
                part = ctx.ik_download()
                assert np.array_equal(part, full[g0:g1]), "shard [%d,%d) of %d differs" % (g0, g1, world)
                kept.append(ctx.counts()["n_live"])
            assert max(kept) < n_all, "no particle was culled: %s of %d" % (kept, n_all)
            if world == 8:
                assert max(kept) < 0.5 * n_all
    # a culled frame cannot serve global sums, and cannot serve another shard
    with pytest.raises(_capi.PgtError):
        ctx.thermo(_capi.OBS_VIRIAL)
    ctx.set_shard(0, G // 8)
    with pytest.raises(_capi.PgtError):
        ctx.ik_compute(_capi.OBS_VIRIAL)
    # slab_cull = 0: every rank builds the whole frame, same bits
    ctx.set_option("slab_cull", 0)
    ctx.set_shard(G // 3, G // 2)
    ctx.upload(fr["position"], fr["diameter"], fr["box"])
    ctx.build_cells()
    ctx.build_bonds()
    ctx.ik_compute(_capi.OBS_VIRIAL)
    assert ctx.counts()["n_live"] == n_all
    assert np.array_equal(ctx.ik_download(), full[G // 3:G // 2])
    ctx.close()


def _rank_body(rank, world, uid, fr, out, errors):
    try:
        import torch
        dev = rank % max(1, _capi.device_count())
        ctx = _capi.Context(dev)
        _setup(ctx, fr)
        ctx.comm_init(world, rank, uid)
        G = NMAX * NMAX
        shards = sharding.grid_shards(G, world)
        N = len(fr["diameter"])
        chunk = (N + world - 1) // world
        first = rank * chunk
        last = min(N, first + chunk)
        # --- the frame: 1/world of it over PCIe, assembled over NVLink
        ctx.set_shard(*shards[rank])
        ctx.upload_part(fr["position"][first:last], fr["diameter"][first:last], first, N, fr["box"])
        ctx.comm_allgather_frame()
        ctx.build_cells()
        ctx.build_bonds()
        ctx.ik_compute(_capi.OBS_VIRIAL)
        mine = ctx.ik_download()
        # --- final field gather on the device
        torch.cuda.set_device(dev)
        gathered = torch.empty((G, 4), dtype=torch.float64, device="cuda:%d" % dev)
        ctx.comm_allgather_field(_capi.FIELD_VIRIAL, [4 * (e - b) for b, e in shards], gathered.data_ptr())
        ctx.sync()
        # --- thermo: whole frame on every rank, pairs summed per particle share, one all-reduce
        ctx.set_shard(0, G)
        ctx.upload_part(fr["position"][first:last], fr["diameter"][first:last], first, N, fr["box"])
        ctx.comm_allgather_frame()
        ctx.build_cells()
        ctx.thermo_set_shard(rank, world)
        t = ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY)
        vec = np.concatenate([t["virial"], [t["energy"]]])
        red = ctx.comm_allreduce_sum(vec)
        out[rank] = dict(mine=mine, gathered=gathered.cpu().numpy(), thermo=red, shard=shards[rank])
        ctx.close()
    except Exception as e:                                   # surfaced by the main thread
        errors.append((rank, repr(e)))


def _run_ranks(world):
    fr = _frame()
    uid = _capi.Context.comm_unique_id()
    out, errors = {}, []
    threads = [threading.Thread(target=_rank_body, args=(r, world, uid, fr, out, errors)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors
    assert len(out) == world
    full, thermo, _ = _full_field(fr)
    want = np.concatenate([thermo["virial"], [thermo["energy"]]])
    for r in range(world):
        b, e = out[r]["shard"]
        assert np.array_equal(out[r]["mine"], full[b:e])
        assert np.array_equal(out[r]["gathered"], full), "rank %d: gathered field differs from the single-GPU field" % r
        assert np.all(np.abs(out[r]["thermo"] - want) <= 1e-13 * np.abs(want) + 1e-300), (out[r]["thermo"], want)


def test_nccl_entry_points_with_one_rank():
    """pgt_frame_upload_part / pgt_comm_allgather_frame / pgt_comm_allgather_field / pgt_comm_allreduce_sum as a
    one-rank communicator: runs on every GPU box."""
    _run_ranks(1)


def test_two_gpus_grid_shard_field_gather_and_thermo_reduce():
    if _capi.device_count() < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")
    _run_ranks(2)
