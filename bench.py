#!/usr/bin/env python
"""bench.py -- IK stress-field throughput (grid-point*frames/s) on 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg5|cfg3|cfg4|cfg1s]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one full pass of the hot path (stages 1-5: cell list, pair forces + half-bond list, IK
gather) over ONE synthetic trajectory frame per GPU.  Frames are sharded over the ranks (each rank
owns different frames, every GPU holds its whole frame: weak scaling, no data-path collective);
`--shard grid` instead splits the grid points of the same frame over the ranks (np.array_split
boundaries, like the reference's MPI ranks) and all-gathers the field with NCCL.

Printed JSON line (rank 0):
  value     grid-point*frames/s with the frames already resident in HBM (CUDA events on the ctx stream,
            max over ranks)
  e2e       same metric through the C-ABI with HOST buffers: pinned H2D of the float32 frame, the
            stages, D2H of the FP64 stress field, every step
  roofline  the IK gather kernel against the FP64 FMA pipe (the path is not a dense contraction and the
            gather is not HBM-bound): achieved = sum_x U(x)*(15q+9) flops / launch time
  cpu_baseline  the oracle (oracle/, the reference's algorithm restated) on the host cores, bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "IK stress-field grid-point*frames/sec (2D polydisperse12)"
UNIT = "grid-point*frames/s"

WORKLOADS = {
    # name: (n = sqrt(N), nmax, potential, cg kernel, cg rcut, GL order, seed key, description)
    "cfg5": dict(n=2048, nmax=2048, pot="poly12", cg=("octic", 2.0, 5), seed="cfg5",
                 desc="BASELINE configs[4]: 2D polydisperse12, synthetic N=4194304 at rho=1, 2048^2 grid, octic R=2 GL-5"),
    "cfg3": dict(n=1024, nmax=1024, pot="poly12", cg=("octic", 2.0, 5), seed="cfg3",
                 desc="BASELINE configs[2]: 2D polydisperse12, synthetic N=1048576 at rho=1, 1024^2 grid, octic R=2 GL-5"),
    "cfg4": dict(n=1024, nmax=1024, pot="poly10", cg=("octic", 4.0, 5), seed="cfg4",
                 desc="BASELINE configs[3]: polydisperse10 (q=3), octic R=4 GL-5, N=1048576, 1024^2 grid"),
    "cfg1s": dict(n=32, nmax=127, pot="poly12", cg=("octic", 2.0, 5), seed="cfg3",
                  desc="small smoke workload (N=1024, 127^2 grid)"),
}


def pot_params(name):
    """Coefficients as the reference's Python factories compute them (potential/__init__.py:78-143)."""
    if name == "poly12":
        v0, eps, rc = 1.0, 0.2, 1.25
        return "poly12", rc, [v0, eps, -28.0 * v0 / rc**12, 48.0 * v0 / rc**14, -21.0 * v0 / rc**16]
    if name == "poly10":
        v0, eps, rc = 1.0, 0.0416667, 1.48
        return "poly10", rc, [v0, eps, -(56.0) * v0 / (rc**10), (140.0) * v0 / (rc**12), -(120.0) * v0 / (rc**14),
                              (35.0) * v0 / (rc**16)]
    raise KeyError(name)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, windows):
        sm, smax, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.lines:
            if not any(a <= t <= b for a, b in windows):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                smax = max(smax, float(f[1]))
                for nm, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        if not sm:   # region shorter than the sampling period: fall back to all samples
            for t, line in self.lines:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[0]))
                    smax = max(smax, float(f[1]))
                except Exception:
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_frames(wl, rank, nframes):
    from pyglasstools_b200 import synth
    frames = []
    for k in range(nframes):
        seed = synth.SEEDS[wl["seed"]] + rank * 1000 + k
        frames.append(synth.synth_frame(wl["n"], seed))
    return frames


def _parse_cpulist(txt):
    cpus = set()
    for part in txt.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def bind_to_gpu_numa_node(local_rank):
    """One process per GPU: keep this rank's host threads, and therefore the pinned staging buffers they first touch,
    on the NUMA node the GPU hangs off (sysfs numa_node of its PCI function).  Returns a description for the log."""
    if os.environ.get("PGT_BENCH_NUMA", "1") == "0":
        return "off"
    try:
        import torch
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        if node < 0:
            return "gpu %s: no NUMA node reported" % bdf
        cpus = _parse_cpulist(open("/sys/devices/system/node/node%d/cpulist" % node).read())
        mine = cpus & os.sched_getaffinity(0)
        if not mine:
            return "gpu %s on node %d: none of its cpus in this process's affinity" % (bdf, node)
        os.sched_setaffinity(0, mine)
        return "gpu %s -> node %d, %d cpus" % (bdf, node, len(mine))
    except Exception as e:                      # sysfs layout differs or the call is not permitted: run unbound
        return "unbound (%s)" % e


def run_reference(args, wl, rank, world):
    """The reference's algorithm on the host cores: the oracle (the reference itself cannot be built
    here -- no Boost / GSL / MPI), all threads, a bounded sample of the grid per step."""
    if rank != 0:
        return
    from oracle import orc
    from pyglasstools_b200 import synth
    # torchrun exports OMP_NUM_THREADS=1 when it starts more than one rank; this arm is rank 0 alone on the whole host
    orc.set_num_threads(len(os.sched_getaffinity(0)))
    fr = synth.synth_frame(wl["n"], synth.SEEDS[wl["seed"]])
    name, rc, prm = pot_params(wl["pot"])
    P = orc.potential(name)
    cg = orc.cgfunc(*wl["cg"])
    t0 = time.time()
    S = orc.System(fr["box"], fr["position"], fr["diameter"])
    build_s = time.time() - t0
    L = float(fr["box"][0])
    axis = synth.linspace_axis(L, wl["nmax"])
    stride = max(1, int(round(wl["nmax"] / 512.0)))
    ii = np.arange(0, wl["nmax"], stride)
    grid = np.zeros((len(ii) ** 2, 3))
    grid[:, 0] = np.repeat(axis[ii], len(ii))
    grid[:, 1] = np.tile(axis[ii], len(ii))
    mr = S.max_rcut(P)
    for _ in range(max(args.warmup, 1)):
        S.ik(P, cg, grid[: len(grid) // 8], max_rcut=mr)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        S.ik(P, cg, grid, max_rcut=mr)
    dt = (time.perf_counter() - t0) / args.steps
    value = len(grid) / dt
    sample = "%d of %d grid points (every %d-th in i and j) of one frame per step; frame build %.2f s excluded" % (
        len(grid), wl["nmax"] ** 2, stride, build_s)
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": args.workload, "description": wl["desc"], "N": wl["n"] ** 2, "grid": wl["nmax"] ** 2},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": orc.num_threads(), "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-field", default="full", choices=("full", "sym"),
                    help="end-to-end read-back: the reference's G x 4 layout, or the symmetric-packed G x 3 (xx, xy, yy)")
    ap.add_argument("--e2e-threads", type=int, default=2, choices=(1, 2), help="1: one host thread drives every end-to-end lane; 2: one host thread per lane")
    ap.add_argument("--e2e-lanes", type=int, default=3, choices=(2, 3, 4), help="contexts (streams) the end-to-end trajectory loop is pipelined over")
    ap.add_argument("--shard", default="frames", choices=["frames", "grid"])
    ap.add_argument("--pool", type=int, default=3, help="distinct frames cycled per rank")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-stride", type=int, default=0)
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return

    import torch
    import torch.distributed as dist
    from pyglasstools_b200 import _capi, sharding, synth

    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else "single rank: unbound"
    if os.environ.get("PGT_BENCH_VERBOSE"):
        print("rank %d: %s, affinity %d cpus" % (rank, numa, len(os.sched_getaffinity(0))), file=sys.stderr, flush=True)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    ctx = _capi.Context(local_rank)
    name, rc, prm = pot_params(wl["pot"])
    ctx.set_potential(name, rc, prm)
    ctx.set_cg(*wl["cg"])
    n, nmax = wl["n"], wl["nmax"]
    N, G = n * n, nmax * nmax
    q = wl["cg"][2]

    # ---- inputs: a small pool of distinct frames per rank, in pinned host memory and resident in HBM
    pool = max(1, args.pool)
    frames = make_frames(wl, rank if args.shard == "frames" else 0, pool)
    L = float(frames[0]["box"][0])
    dmax = max(float(f["diameter"].max()) for f in frames)
    ctx.set_max_rcut(max(rc * dmax, dmax))                   # Calculator.h:37-46, frozen
    axis = synth.linspace_axis(L, nmax)
    ctx.set_grid_cartesian(axis, axis)
    if args.shard == "grid" and world > 1:
        shards = sharding.grid_shards(G, world)          # np.array_split boundaries = the reference's MPI ranks
        sizes = [e - b for b, e in shards]
        ctx.set_shard(*shards[rank])
        g_local = sizes[rank]
    else:
        sizes = [G]
        g_local = G
    pin_pos = [_capi.PinnedArray((N, 3), np.float32) for _ in range(pool)]
    pin_diam = [_capi.PinnedArray((N,), np.float32) for _ in range(pool)]
    dev_pos, dev_diam = [], []
    for k, f in enumerate(frames):
        pin_pos[k].array[...] = f["position"]
        pin_diam[k].array[...] = f["diameter"]
        dev_pos.append(torch.from_numpy(f["position"]).cuda())
        dev_diam.append(torch.from_numpy(f["diameter"]).cuda())
    ncomp_out = 3 if args.e2e_field == "sym" else 4
    pin_out = _capi.PinnedArray((g_local, ncomp_out), np.float64)
    box = frames[0]["box"]
    torch.cuda.synchronize()

    def step_resident(k):
        ctx.set_device_frame(dev_pos[k].data_ptr(), dev_diam[k].data_ptr(), N, box, is_f32=True)
        ctx.build_cells()
        ctx.build_bonds()
        ctx.ik_compute(_capi.OBS_VIRIAL)

    # end to end: the trajectory loop pipelined over --e2e-lanes contexts (one stream each): frame s+1's upload and
    # stages overlap frame s's field read-back; every frame is uploaded from and read back to pinned host memory
    lanes = [(ctx, pin_out)]
    for _ in range(args.e2e_lanes - 1):
        ctx_b = _capi.Context(local_rank)
        ctx_b.set_potential(name, rc, prm)
        ctx_b.set_cg(*wl["cg"])
        ctx_b.set_max_rcut(max(rc * dmax, dmax))
        ctx_b.set_grid_cartesian(axis, axis)
        if args.shard == "grid" and world > 1:
            ctx_b.set_shard(*shards[rank])
        lanes.append((ctx_b, _capi.PinnedArray((g_local, ncomp_out), np.float64)))
    nl = len(lanes)

    def step_e2e(s, k):
        c, out = lanes[s % nl]
        c.upload(pin_pos[k].array, pin_diam[k].array, box)
        c.build_cells()
        c.build_bonds()
        c.ik_compute(_capi.OBS_VIRIAL)
        c.ik_download_wait()                       # this lane's previous field has landed in `out` (consumed here)
        if args.e2e_field == "sym":
            c.ik_download_sym_async(_capi.FIELD_VIRIAL, out.array)
        else:
            c.ik_download_async(_capi.FIELD_VIRIAL, out.array)

    # ---- algorithmic work of the gather per frame (outside the timed regions)
    flops_per_frame, items_per_frame, hb_per_frame = [], [], []
    for k in range(pool):
        step_resident(k)
        w = ctx.ik_count_work()
        flops_per_frame.append(w["unique_bonds"] * (15 * q + 9))
        items_per_frame.append(w["halfbond_items"])
        hb_per_frame.append(ctx.counts()["n_halfbonds"])
    fp64_peak = ctx.fp64_peak()                               # DFMA-loop probe, TFLOP/s, this GPU, this run

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- device-resident timing -------------------------------------------------------------
    for s in range(args.warmup):
        step_resident(s % pool)
    barrier()
    ctx.sync()
    launches0 = ctx.launches()
    stage_ms = {"cells": [], "bonds": [], "ik": []}
    w0 = time.time()
    ctx.event_record(0)
    for s in range(args.steps):
        step_resident(s % pool)
        # build_cells/build_bonds synchronise the stream, so the previous stages' events are complete
        stage_ms["cells"].append(ctx.stage_ms("cells"))
        stage_ms["bonds"].append(ctx.stage_ms("bonds"))
    ctx.event_record(1)
    ctx.sync()
    barrier()
    w1 = time.time()
    launches_resident = ctx.launches() - launches0
    ms_resident = max_over_ranks(ctx.event_elapsed_ms(0, 1))

    # the gather kernel alone, timed live over the same region (separate pass so that reading the
    # event does not perturb the step loop above)
    ik_ms = []
    for s in range(min(args.steps, 50)):
        step_resident(s % pool)
        ik_ms.append((s % pool, ctx.stage_ms("ik")))

    # ---- end to end: host buffers in, host field out -----------------------------------------
    for s in range(max(args.warmup, 3)):
        step_e2e(s, s % pool)
    for c, _ in lanes:
        c.sync()
    barrier()
    w2 = time.time()
    ctx.event_record(2)
    if args.e2e_threads == 2:
        # one host thread per lane (ctypes releases the GIL inside the C-ABI calls): while one lane's host thread waits
        # for a stage's counts, the other keeps the device fed
        import threading

        def lane_worker(lane):
            for s in range(lane, args.steps, nl):
                step_e2e(s, s % pool)

        workers = [threading.Thread(target=lane_worker, args=(lane,)) for lane in range(nl)]
        for t in workers:
            t.start()
        for t in workers:
            t.join()
    else:
        for s in range(args.steps):
            step_e2e(s, s % pool)
    for c, _ in reversed(lanes):                   # every lane drained: the end event below is later than all work
        c.sync()
    ctx.event_record(3)
    ctx.sync()
    barrier()
    w3 = time.time()
    ms_e2e = max_over_ranks(ctx.event_elapsed_ms(2, 3))
    if rank == 0:
        sampler.stop()

    if args.shard == "grid" and world > 1:
        # final field gather over NVLink (one NCCL broadcast per shard inside a group), checked once
        uid = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(world, rank, uid[0])
        full = torch.empty((G, 4), dtype=torch.float64, device="cuda")
        ctx.comm_allgather_field(_capi.FIELD_VIRIAL, [4 * s for s in sizes], full.data_ptr())
        ctx.sync()

    # ---- report ------------------------------------------------------------------------------
    units_per_step = G if (args.shard == "frames" or world == 1) else G / world
    total_units = sum_over_ranks(units_per_step * args.steps)
    value = total_units / (ms_resident * 1e-3)
    e2e_value = total_units / (ms_e2e * 1e-3)
    launches_total = int(sum_over_ranks(launches_resident))
    e2e_check = float(sum(np.abs(o.array).sum() for _, o in lanes))   # the read-backs are real
    assert np.isfinite(e2e_check) and e2e_check > 0
    ik_avg_ms = float(np.mean([m for _, m in ik_ms]))
    flops_avg = float(np.mean([flops_per_frame[k] for k, _ in ik_ms])) * (g_local / G if args.shard == "grid" else 1.0)
    achieved = flops_avg / (ik_avg_ms * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    # DRAM bytes of one gather launch from the committed `ncu --set full` capture of this workload (profiles/)
    traffic = None
    try:
        if args.workload == "cfg5" and args.shard == "frames":
            for line in open(os.path.join(ROOT, "profiles", "r1_ik_gather2_full_cfg5.txt")):
                if line.startswith("traffic"):
                    traffic = float(line.split()[-1])
    except Exception:
        pass
    h2d = N * 3 * 4 + N * 4
    d2h = g_local * ncomp_out * 8

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_resident / args.steps, "higher_is_better": True,
            "scaling": "weak" if args.shard == "frames" else "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": args.workload, "description": wl["desc"], "N": N, "grid": G, "potential": wl["pot"],
                       "cg": {"kernel": wl["cg"][0], "rcut": wl["cg"][1], "order": q}, "shard": args.shard,
                       "frames_per_step_per_gpu": 1 if args.shard == "frames" else 1.0 / world, "frame_pool": pool,
                       "l2": "inputs larger than L2: each step rebuilds and streams a %.0f MB half-bond list + %.0f MB sorted frame"
                             % (np.mean(hb_per_frame) * 36 / 1e6, N * 28 / 1e6),
                       "stages_timed": "1-5 (cells, pair forces + half-bonds, IK gather, virial field)"},
            "clocks": sampler.summary([(w0, w1), (w2, w3)]),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "host_binding": numa, "api": "C-ABI pgt_frame_upload .. %s, pinned host buffers, frames pipelined over %d contexts, %s" % ("pgt_ik_download_sym_async (xx,xy,yy)" if args.e2e_field == "sym" else "pgt_ik_download_async", args.e2e_lanes, "one host thread per context" if args.e2e_threads == 2 else "one host thread")},
            "gpu_launches": launches_total,
            "stage_ms": {"cells": float(np.mean(stage_ms["cells"])), "bonds": float(np.mean(stage_ms["bonds"])), "ik": ik_avg_ms},
            "roofline": {"kernel": "k_ik_gather", "bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp64_peak,
                         "peak_source": "DFMA-loop probe run on this GPU in this process (MEASURED_PEAKS.json has no FP64 entry); "
                                        "nominal 148 SM x 64 FMA/clk x 2 x 1.965 GHz = 37.2",
                         "frac_of_nominal_37.2": achieved / 37.2,
                         "work": "sum_x U(x)*(15q+9) flops, U counted on device (pgt_ik_count_work), %.4g flops per launch"
                                 % flops_avg,
                         "items_per_unique_bond": float(np.mean(items_per_frame)) / max(1.0, float(np.mean(flops_per_frame)) / (15 * q + 9)),
                         "traffic": traffic,
                         "hbm_peak_gbs": peaks.get("hbm_gbs")},
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(wl, frames[0], args)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(wl, fr, args):
    """The oracle on the box's host cores, on a bounded strided sample of the same frame's grid."""
    from oracle import orc
    from pyglasstools_b200 import synth
    name, rc, prm = pot_params(wl["pot"])
    P = orc.potential(name)
    cg = orc.cgfunc(*wl["cg"])
    t0 = time.time()
    S = orc.System(fr["box"], fr["position"], fr["diameter"])
    build_s = time.time() - t0
    L = float(fr["box"][0])
    nmax = wl["nmax"]
    axis = synth.linspace_axis(L, nmax)
    stride = args.cpu_sample_stride or 1
    ii = np.arange(0, nmax, stride)
    grid = np.zeros((len(ii) ** 2, 3))
    grid[:, 0] = np.repeat(axis[ii], len(ii))
    grid[:, 1] = np.tile(axis[ii], len(ii))
    mr = S.max_rcut(P)
    S.ik(P, cg, grid[:4096], max_rcut=mr)
    passes, t0 = 0, time.perf_counter()
    while passes < 1 or (time.perf_counter() - t0 < 10.0 and passes < 64):     # >= 10 s of CPU work
        S.ik(P, cg, grid, max_rcut=mr)
        passes += 1
    dt = (time.perf_counter() - t0) / passes
    return {"value": len(grid) / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": "%d of %d grid points (every %d-th in i and j) of one frame, %d passes of %.1f s; oracle frame build %.1f s excluded"
                      % (len(grid), nmax * nmax, stride, passes, dt, build_s)}


if __name__ == "__main__":
    main()
