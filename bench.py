#!/usr/bin/env python
"""bench.py -- IK stress-field throughput (grid-point*frames/s) on 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg5|cfg3|cfg4|cfg1s]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one full pass of the hot path (stages 1-5: cell list, pair forces + half-bond list, IK
gather) over ONE synthetic trajectory frame per GPU.  Frames are sharded over the ranks (each rank
owns different frames, every GPU holds its whole frame: weak scaling, no data-path collective);
`--shard grid` instead splits the grid points of the same frame over the ranks (np.array_split
boundaries, like the reference's MPI ranks) and all-gathers the field with NCCL.

At N > 1 the same run also times ONE frame split over the ranks by grid points (`grid_shard` in the JSON line): parts
of the frame pushed over PCIe and assembled with ncclAllGather, slab-local cell / bond lists, every rank reading back
its own shard; the gathered field is compared bit for bit with the single-GPU field once, outside the timing.

Printed JSON line (rank 0):
  value     grid-point*frames/s with the frames already resident in HBM (CUDA events on the ctx stream,
            max over ranks)
  e2e       same metric through the C-ABI with HOST buffers: pinned H2D of the float32 frame, the
            stages, D2H of the FP64 stress field, every step
  roofline  the IK gather kernel against the FP64 FMA pipe (the path is not a dense contraction and the
            gather is not HBM-bound): achieved = sum_x U(x)*(15q+9) flops / launch time
  roofline_stages  cell-list and bond stages against the measured HBM copy bandwidth (MEASURED_PEAKS.json)
  other_configs    BASELINE configs 1-4 (shipped fixture, thermo trajectory, cfg3, cfg4), short runs in the same process
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


def config_dict(args, wl, world, pool):
    """The workload description both arms print (the driver compares them)."""
    n, nmax, q = wl["n"], wl["nmax"], wl["cg"][2]
    return {"workload": args.workload, "description": wl["desc"], "N": n * n, "grid": nmax * nmax, "potential": wl["pot"],
            "cg": {"kernel": wl["cg"][0], "rcut": wl["cg"][1], "order": q}, "shard": args.shard,
            "frames_per_step_per_gpu": 1 if args.shard == "frames" else 1.0 / world, "frame_pool": pool,
            "l2": "inputs larger than L2: each step rebuilds and streams a ~%.0f MB half-bond list + %.0f MB sorted frame"
                  % (n * n * 3.61 * 36 / 1e6, n * n * 28 / 1e6),
            "stages_timed": "1-5 (cells, pair forces + half-bonds, IK gather, virial field)"}


def oracle_rate(wl, fr, stride, min_seconds, warm):
    """grid points / s of the oracle's IK loop (all host threads, timing build) on every stride-th grid row and column."""
    os.environ["PGT_ORACLE_FAST"] = "1"                  # before the first oracle call: -O3 -march=native copy
    from oracle import orc
    from pyglasstools_b200 import synth
    orc.set_num_threads(len(os.sched_getaffinity(0)))    # torchrun exports OMP_NUM_THREADS=1 for N > 1
    name, rc, prm = pot_params(wl["pot"])
    P = orc.potential(name)
    cg = orc.cgfunc(*wl["cg"])
    t0 = time.time()
    S = orc.System(fr["box"], fr["position"], fr["diameter"])
    build_s = time.time() - t0
    L = float(fr["box"][0])
    nmax = wl["nmax"]
    axis = synth.linspace_axis(L, nmax)
    ii = np.arange(0, nmax, stride)
    grid = np.zeros((len(ii) ** 2, 3))
    grid[:, 0] = np.repeat(axis[ii], len(ii))
    grid[:, 1] = np.tile(axis[ii], len(ii))
    mr = S.max_rcut(P)
    for _ in range(warm):
        S.ik(P, cg, grid[: max(4096, len(grid) // 8)], max_rcut=mr)
    passes, t0 = 0, time.perf_counter()
    while passes < 1 or (time.perf_counter() - t0 < min_seconds and passes < 64):
        S.ik(P, cg, grid, max_rcut=mr)
        passes += 1
    dt = (time.perf_counter() - t0) / passes
    return dict(rate=len(grid) / dt, dt=dt, passes=passes, points=len(grid), stride=stride, build_s=build_s,
                cores=orc.num_threads(), lib=getattr(orc.lib(), "_pgt_build", "liborc.so"))


def run_reference(args, wl, rank, world):
    """The reference's algorithm on the host cores: the oracle (the reference itself cannot be built
    here -- no Boost / GSL / MPI), all threads, a bounded sample of the grid per step."""
    if rank != 0:
        return
    from pyglasstools_b200 import synth
    fr = synth.synth_frame(wl["n"], synth.SEEDS[wl["seed"]])
    stride = args.cpu_sample_stride or max(1, int(round(wl["nmax"] / 512.0)))
    # K timed steps, each one pass over the bounded sample
    os.environ["PGT_ORACLE_FAST"] = "1"
    from oracle import orc
    orc.set_num_threads(len(os.sched_getaffinity(0)))
    name, rc, prm = pot_params(wl["pot"])
    P = orc.potential(name)
    cg = orc.cgfunc(*wl["cg"])
    t0 = time.time()
    S = orc.System(fr["box"], fr["position"], fr["diameter"])
    build_s = time.time() - t0
    L = float(fr["box"][0])
    axis = synth.linspace_axis(L, wl["nmax"])
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
    sample = "%d of %d grid points (every %d-th in i and j) of one frame per step, rate extrapolated to the grid; frame build %.2f s excluded; %s (-O3 -march=native -fopenmp, the reference's flags)" % (
        len(grid), wl["nmax"] ** 2, stride, build_s, getattr(orc.lib(), "_pgt_build", "liborc.so"))
    cfg = config_dict(args, wl, 1 if args.shard == "frames" else world, max(1, args.pool))
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
           "scaling": "weak" if args.shard == "frames" else "strong",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": orc.num_threads(), "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def setup_ctx(_capi, synth, device, wl, dmax, L):
    ctx = _capi.Context(device)
    name, rc, prm = pot_params(wl["pot"])
    ctx.set_potential(name, rc, prm)
    ctx.set_cg(*wl["cg"])
    ctx.set_max_rcut(max(rc * dmax, dmax))                   # Calculator.h:37-46, frozen
    axis = synth.linspace_axis(L, wl["nmax"])
    ctx.set_grid_cartesian(axis, axis)
    return ctx


LANE_ASYNC = int(os.environ.get("PGT_BENCH_LANE_ASYNC", "0"))   # developer knob: deferred counts inside multi-context e2e lanes


def stage_bytes(N, ncells, n_hb):
    """Algorithmic HBM bytes per frame of the two memory-bound stages:
    cells  60 B / particle (16 raw float32 read + 4 key + 4 slot + 4 sorted id + 32 sorted x, y, (x,y), d written)
           + 8 B / cell (count written, start read);
    bonds  24 B / particle read (x, y, d) + 8 B / particle (offsets) + 36 B / half-bond record written."""
    return {"cells": 60.0 * N + 8.0 * ncells, "bonds": 32.0 * N + 36.0 * n_hb}


def quick_workload(_capi, synth, device, name, steps=12, warm=3):
    """A short device-timed run of another BASELINE configuration (same stages, same timing rules): for `other_configs`."""
    wl = WORKLOADS[name]
    fr = synth.synth_frame(wl["n"], synth.SEEDS[wl["seed"]])
    import torch
    ctx = setup_ctx(_capi, synth, device, wl, float(fr["diameter"].max()), float(fr["box"][0]))
    dpos, ddiam = torch.from_numpy(fr["position"]).cuda(), torch.from_numpy(fr["diameter"]).cuda()
    N, G, q = wl["n"] ** 2, wl["nmax"] ** 2, wl["cg"][2]

    def step():
        ctx.set_device_frame(dpos.data_ptr(), ddiam.data_ptr(), N, fr["box"], is_f32=True)
        ctx.build_cells()
        ctx.build_bonds()
        ctx.ik_compute(_capi.OBS_VIRIAL)
    for _ in range(warm):
        step()
    w = ctx.ik_count_work()
    ctx.sync()
    ctx.event_record(0)
    for _ in range(steps):
        step()
    ctx.event_record(1)
    ctx.sync()
    ms = ctx.event_elapsed_ms(0, 1) / steps
    ik = ctx.stage_ms("ik")
    flops = w["unique_bonds"] * (15 * q + 9)
    out = {"description": wl["desc"], "value": G / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
           "stage_ms": {"cells": ctx.stage_ms("cells"), "bonds": ctx.stage_ms("bonds"), "ik": ik},
           "ik_tflops": flops / (ik * 1e-3) / 1e12, "ik_frac_of_nominal_37.2": flops / (ik * 1e-3) / 1e12 / 37.2, "steps": steps}
    ctx.close()
    return out


def fixture_workload(_capi, synth, device):
    """BASELINE configs[0]: the shipped example (example_scripts/irvingkirkwood/test-irvingkirkwood-module.py:11-25 on
    polydisperse12.gsd: poly12, octic R = 2 GL-5, dx = 0.25, frames 180..199), frames from the committed fixture, through the
    C-ABI with host buffers (upload + stages + field read-back per frame)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "polydisperse12_frames.npz"))
    diam, box = z["diameter"], z["box"]
    L = float(box[0])
    nmax = int(L / 0.25)                                      # ikcalculator: nmax = int(L / dx), linspace both ends
    frames = []
    for f in range(180, 200):
        p = np.zeros((z["xy"].shape[1], 3), np.float32)
        p[:, :2] = z["xy"][f]
        frames.append(p)
    ctx = _capi.Context(device)
    name, rc, prm = pot_params("poly12")
    ctx.set_potential(name, rc, prm)
    ctx.set_cg("octic", 2.0, 5)
    dmax = float(diam.max())
    ctx.set_max_rcut(max(rc * dmax, dmax))
    axis = synth.linspace_axis(L, nmax)
    ctx.set_grid_cartesian(axis, axis)
    out = np.empty((nmax * nmax, 4))
    pin = [_capi.PinnedArray(p.shape, np.float32) for p in frames]
    for a, p in zip(pin, frames):
        a.array[...] = p
    pdiam = _capi.PinnedArray(diam.shape, np.float32)
    pdiam.array[...] = diam
    outs = [_capi.PinnedArray((nmax * nmax, 4), np.float64) for _ in range(4)]

    def run_sync():
        for p in frames:
            ctx.upload(p, diam, box)
            ctx.build_cells()
            ctx.build_bonds()
            ctx.ik_compute(_capi.OBS_VIRIAL)
            ctx.ik_download(_capi.FIELD_VIRIAL, out)

    def run():
        # page-locked buffers, counts deferred, read-backs four deep: the host never waits for the frame it has just submitted
        for s, a in enumerate(pin):
            ctx.upload(a.array, pdiam.array, box)
            ctx.build_cells()
            ctx.build_bonds()
            ctx.ik_compute(_capi.OBS_VIRIAL)
            ctx.ik_download_async(_capi.FIELD_VIRIAL, outs[s % 4].array)
            ctx.ik_download_wait_keep(3)
        ctx.ik_download_wait()
        ctx.sync()
    run_sync()
    ref_last = out.copy()
    run()
    assert np.array_equal(outs[(len(pin) - 1) % 4].array, ref_last), "pipelined field differs from the synchronous one"
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        run()
    dt = (time.perf_counter() - t0) / (reps * len(frames))
    t0 = time.perf_counter()
    for _ in range(reps):
        run_sync()
    dt_sync = (time.perf_counter() - t0) / (reps * len(frames))
    ctx.close()
    return {"description": "BASELINE configs[0]: shipped polydisperse12.gsd frames 180..199 (N = %d), script's own grid %d^2 (dx = 0.25), poly12, octic R=2 GL-5; through the C-ABI from page-locked host buffers, field read back every frame (pgt_ik_download_async four deep, counts deferred)"
                           % (len(diam), nmax), "value": nmax * nmax / dt, "unit": UNIT, "ms_per_frame": dt * 1e3, "frames": len(frames),
            "synchronous": {"value": nmax * nmax / dt_sync, "ms_per_frame": dt_sync * 1e3, "api": "pageable buffers, pgt_ik_download (waits) every frame"}}


def thermo_workload(_capi, synth, device, nframes=1000):
    """BASELINE configs[1]: thermo global virial stress + potential energy, synthetic N = 65536, a 1000-frame trajectory
    streamed from host memory through the C-ABI (upload + cell list + pair sums + 25-double read-back per frame)."""
    pool = [synth.synth_frame(256, synth.SEEDS["cfg2"] + f) for f in range(8)]
    ctx = _capi.Context(device)
    name, rc, prm = pot_params("poly12")
    ctx.set_potential(name, rc, prm)
    dmax = max(float(fr["diameter"].max()) for fr in pool)
    ctx.set_max_rcut(max(rc * dmax, dmax))
    mask = _capi.OBS_VIRIAL | _capi.OBS_ENERGY
    # page-locked frames and results; pgt_thermo_compute_async + deferred counts: the loop never waits for the device, the
    # host collects the results of 32 frames at a time
    N = len(pool[0]["diameter"])
    pin_pos = [_capi.PinnedArray((N, 3), np.float32) for _ in pool]
    pin_diam = [_capi.PinnedArray((N,), np.float32) for _ in pool]
    for k, fr in enumerate(pool):
        pin_pos[k].array[...] = fr["position"]
        pin_diam[k].array[...] = fr["diameter"]
    ring = [_capi.PinnedArray((25,), np.float64) for _ in range(32)]
    box = pool[0]["box"]
    check = []
    for f in range(16):
        k = f % len(pool)
        ctx.upload(pin_pos[k].array, pin_diam[k].array, box)
        ctx.build_cells()
        t = ctx.thermo(mask)
        if f < len(pool):
            check.append(np.concatenate([t["virial"], t["kinetic"], [t["energy"]], t["born"]]))
    t0 = time.perf_counter()
    for f in range(nframes):
        k = f % len(pool)
        ctx.upload(pin_pos[k].array, pin_diam[k].array, box)
        ctx.build_cells()
        ctx.thermo_async(mask, ring[f % len(ring)].array)
        if f % len(ring) == len(ring) - 1:
            ctx.sync()                                   # the 32 results are consumed here
    ctx.sync()
    dt = time.perf_counter() - t0
    last = (nframes - 1) % len(pool)
    assert np.array_equal(ring[(nframes - 1) % len(ring)].array, check[last]), "asynchronous thermo sums differ from the synchronous ones"
    t1 = time.perf_counter()
    for f in range(min(nframes, 200)):
        k = f % len(pool)
        ctx.upload(pin_pos[k].array, pin_diam[k].array, box)
        ctx.build_cells()
        ctx.thermo(mask)
    dt_sync = (time.perf_counter() - t1) / min(nframes, 200)
    out = {"description": "BASELINE configs[1]: thermo virial + energy, synthetic N=65536, %d frames through the C-ABI from page-locked host buffers (pgt_frame_upload, pgt_build_cells, pgt_thermo_compute_async; the host collects 32 frames' results at a time)" % nframes,
           "value": nframes / dt, "unit": "frames/s", "synchronous_frames_per_s": 1.0 / dt_sync,
           "device_ms": {"cells": ctx.stage_ms("cells"), "thermo": ctx.stage_ms("thermo")}}
    ctx.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-field", default="sym", choices=("auto", "full", "sym"),
                    help="end-to-end read-back: the reference's G x 4 layout, or the symmetric-packed G x 3 (xx, xy, yy); auto: full on one GPU, sym on several (the host's PCIe / memory path is shared)")
    ap.add_argument("--e2e-threads", type=int, default=2, choices=(1, 2), help="1: one host thread drives every end-to-end lane; 2: one host thread per lane")
    ap.add_argument("--e2e-lanes", type=int, default=0, choices=(0, 1, 2, 3, 4), help="contexts the end-to-end trajectory loop is spread over. 0 (default): 2 on one GPU (one host thread each: measured fastest, 2.36-2.41 ms per frame at cfg5; the single-context pipeline, 2.54 ms, is timed beside it as e2e_one_context) and 1 under torchrun (N > 1: the host path is the bound there and more host threads per rank do not help); 1: one context and one host thread pipeline upload / stages / read-back on the context's three streams")
    ap.add_argument("--e2e-buffers", type=int, default=3, choices=(2, 3, 4), help="host field buffers the single-context loop reads back into in turn")
    ap.add_argument("--e2e-pos", default="xy", choices=("xy", "xyz"), help="host position rows of the end-to-end loop: xy = N x 2 float32 through pgt_frame_upload_xy (a 2-D frame has no z column: 12 B per particle cross the link; default), xyz = N x 3 through pgt_frame_upload (the GSD layout, 16 B)")
    ap.add_argument("--shard", default="frames", choices=["frames", "grid"])
    ap.add_argument("--pool", type=int, default=3, help="distinct frames cycled per rank")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    ap.add_argument("--no-grid-shard", action="store_true", help="N > 1: skip the grid-sharded measurement of one frame")
    ap.add_argument("--cpu-sample-stride", type=int, default=0)
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.e2e_lanes == 0:
        args.e2e_lanes = 2 if world == 1 else 1
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return
    if args.shard == "grid" and world == 1:
        args.shard = "frames"
    if args.e2e_field == "auto":
        args.e2e_field = "sym"

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

    name, rc, prm = pot_params(wl["pot"])
    n, nmax = wl["n"], wl["nmax"]
    N, G = n * n, nmax * nmax
    q = wl["cg"][2]

    # ---- inputs: a small pool of distinct frames per rank, in pinned host memory and resident in HBM
    pool = max(1, args.pool)
    frames = make_frames(wl, rank if args.shard == "frames" else 0, pool)
    L = float(frames[0]["box"][0])
    dmax = max(float(f["diameter"].max()) for f in frames)
    grid_mode = args.shard == "grid"
    ctx = setup_ctx(_capi, synth, local_rank, wl, dmax, L)
    shards = sharding.grid_shards(G, world)                  # np.array_split boundaries = the reference's MPI ranks
    if grid_mode:
        ctx.set_shard(*shards[rank])
        g_local = shards[rank][1] - shards[rank][0]
    else:
        g_local = G
    pin_pos = [_capi.PinnedArray((N, 3), np.float32) for _ in range(pool)]
    pin_diam = [_capi.PinnedArray((N,), np.float32) for _ in range(pool)]
    pin_xy = [_capi.PinnedArray((N, 2), np.float32) for _ in range(pool)]
    pos_xy = [args.e2e_pos == "xy"]              # the end-to-end loops upload N x 2 rows (the synthetic frames have z = 0)
    dev_pos, dev_diam = [], []
    for k, f in enumerate(frames):
        pin_pos[k].array[...] = f["position"]
        pin_xy[k].array[...] = f["position"][:, :2]
        assert not f["position"][:, 2].any()
        pin_diam[k].array[...] = f["diameter"]
        dev_pos.append(torch.from_numpy(f["position"]).cuda())
        dev_diam.append(torch.from_numpy(f["diameter"]).cuda())
    ncomp_out = 3 if args.e2e_field == "sym" else 4
    box = frames[0]["box"]
    torch.cuda.synchronize()

    comm_ready = [False]

    def comm_init_all(ctxs):
        for c in ctxs:
            uid = [_capi.Context.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            c.comm_init(world, rank, uid[0])

    def step_resident(c, k):
        c.set_device_frame(dev_pos[k].data_ptr(), dev_diam[k].data_ptr(), N, box, is_f32=True)
        c.build_cells()
        c.build_bonds()
        c.ik_compute(_capi.OBS_VIRIAL)

    def make_lanes(first_ctx, nlanes, shard):
        lanes = [(first_ctx, _capi.PinnedArray((shard[1] - shard[0], ncomp_out), np.float64))]
        for _ in range(nlanes - 1):
            cb = setup_ctx(_capi, synth, local_rank, wl, dmax, L)
            cb.set_shard(*shard)
            # several contexts: each stage waits for its count, which keeps the kernels of different frames from
            # co-running on the device (measured: 2.46 against 2.58 ms per frame at cfg5 with deferred counts)
            cb.set_option("async_counts", LANE_ASYNC)
            lanes.append((cb, _capi.PinnedArray((shard[1] - shard[0], ncomp_out), np.float64)))
        return lanes

    chunk = (N + world - 1) // world
    p_first, p_last = rank * chunk, min(N, (rank + 1) * chunk)

    def step_e2e(lanes, s, k, parts):
        c, out = lanes[s % len(lanes)]
        if parts:
            # this rank's 1/world of the frame over PCIe, the rest over NVLink
            c.upload_part(pin_pos[k].array[p_first:p_last], pin_diam[k].array[p_first:p_last], p_first, N, box)
            c.comm_allgather_frame()
        elif pos_xy[0]:
            c.upload_xy(pin_xy[k].array, pin_diam[k].array, box)
        else:
            c.upload(pin_pos[k].array, pin_diam[k].array, box)
        c.build_cells()
        c.build_bonds()
        c.ik_compute(_capi.OBS_VIRIAL)
        c.ik_download_wait()                       # this lane's previous field has landed in `out` (consumed here)
        if args.e2e_field == "sym":
            c.ik_download_sym_async(_capi.FIELD_VIRIAL, out.array)
        else:
            c.ik_download_async(_capi.FIELD_VIRIAL, out.array)

    def time_resident(c, steps, warm):
        for s in range(warm):
            step_resident(c, s % pool)
        barrier()
        c.sync()
        l0 = c.launches()
        t0 = time.time()
        c.event_record(0)
        for s in range(steps):                     # no host synchronisation inside: the frames' counts are deferred
            step_resident(c, s % pool)
        c.event_record(1)
        c.sync()
        barrier()
        t1 = time.time()
        ms = max_over_ranks(c.event_elapsed_ms(0, 1))
        nl = c.launches() - l0
        # per-stage device times, read in a separate pass so that waiting for the events does not perturb the loop above
        st = {"cells": [], "bonds": [], "ik": []}
        for s in range(min(steps, 30)):
            step_resident(c, s % pool)
            for k in st:
                st[k].append(c.stage_ms(k))
        return ms, nl, st, (t0, t1)

    def run_pipelined(c, outs, steps, parts):
        """The trajectory loop in ONE context and one host thread: frame s+1 is uploaded (upload stream) while frame s is
        computed (main stream, counts deferred: no host wait between stages) and frame s-1 is read back (read-back stream)
        into one of a few host buffers used in turn; the host only waits for the oldest read-back in flight."""
        def up(s):
            k = s % pool
            if parts:
                c.upload_part(pin_pos[k].array[p_first:p_last], pin_diam[k].array[p_first:p_last], p_first, N, box)
            elif pos_xy[0]:
                c.upload_xy(pin_xy[k].array, pin_diam[k].array, box)
            else:
                c.upload(pin_pos[k].array, pin_diam[k].array, box)
        up(0)
        for s in range(steps):
            if parts:
                c.comm_allgather_frame()
            c.build_cells()
            c.build_bonds()
            c.ik_compute(_capi.OBS_VIRIAL)
            if s + 1 < steps:
                up(s + 1)
            if args.e2e_field == "sym":
                c.ik_download_sym_async(_capi.FIELD_VIRIAL, outs[s % len(outs)].array)
            else:
                c.ik_download_async(_capi.FIELD_VIRIAL, outs[s % len(outs)].array)
            c.ik_download_wait_keep(len(outs) - 1)   # the oldest of the host buffers has landed (consumed here)
        c.ik_download_wait()

    def time_e2e(lanes, steps, warm, parts, threads):
        nl = len(lanes)
        if nl == 1:
            c = lanes[0][0]
            outs = [lanes[0][1]] + [_capi.PinnedArray(lanes[0][1].array.shape, np.float64) for _ in range(args.e2e_buffers - 1)]
            run_pipelined(c, outs, max(warm, 4), parts)
            c.sync()
            barrier()
            t0 = time.time()
            c.event_record(2)
            run_pipelined(c, outs, steps, parts)
            c.event_record(3)
            c.sync()
            barrier()
            t1 = time.time()
            for o in outs[1:]:
                lanes.append((c, o))               # every buffer takes part in the read-back check
            return max_over_ranks(c.event_elapsed_ms(2, 3)), (t0, t1)
        for s in range(max(warm, 3)):
            step_e2e(lanes, s, s % pool, parts)
        for c, _ in lanes:
            c.sync()
        barrier()
        t0 = time.time()
        lanes[0][0].event_record(2)
        if threads == 2 and nl > 1 and not parts:
            # one host thread per lane (ctypes releases the GIL inside the C-ABI calls): while one lane's host thread waits
            # for a stage's counts, the other keeps the device fed
            def lane_worker(lane):
                for s in range(lane, steps, nl):
                    step_e2e(lanes, s, s % pool, parts)
            workers = [threading.Thread(target=lane_worker, args=(lane,)) for lane in range(nl)]
            for t in workers:
                t.start()
            for t in workers:
                t.join()
        else:
            # collectives must be issued in the same order on every rank: one host thread
            for s in range(steps):
                step_e2e(lanes, s, s % pool, parts)
        for c, _ in reversed(lanes):               # every lane drained: the end event below is later than all work
            c.sync()
        lanes[0][0].event_record(3)
        lanes[0][0].sync()
        barrier()
        t1 = time.time()
        return max_over_ranks(lanes[0][0].event_elapsed_ms(2, 3)), (t0, t1)

    # ---- algorithmic work of the gather per frame (outside the timed regions)
    flops_per_frame, items_per_frame, hb_per_frame, ncells = [], [], [], 0
    for k in range(pool):
        step_resident(ctx, k)
        w = ctx.ik_count_work()
        flops_per_frame.append(w["unique_bonds"] * (15 * q + 9))
        items_per_frame.append(w["halfbond_items"])
        cnt = ctx.counts()
        hb_per_frame.append(cnt["n_halfbonds"])
        ncells = cnt["n_cells"]
    fp64_peak = ctx.fp64_peak()                               # DFMA-loop probe, TFLOP/s, this GPU, this run

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- device-resident timing -------------------------------------------------------------
    ms_resident, launches_resident, stage_ms, win_a = time_resident(ctx, args.steps, args.warmup)

    # the gather kernel alone, timed live over the same region (separate pass so that reading the
    # event does not perturb the step loop above)
    ik_ms = []
    for s in range(min(args.steps, 50)):
        step_resident(ctx, s % pool)
        ik_ms.append((s % pool, ctx.stage_ms("ik")))

    # ---- end to end: host buffers in, host field out -----------------------------------------
    my_shard = shards[rank] if grid_mode else (0, G)
    lanes = make_lanes(ctx, args.e2e_lanes, my_shard)
    if grid_mode:
        comm_init_all([c for c, _ in lanes])
    if len(lanes) > 1:
        ctx.set_option("async_counts", LANE_ASYNC)
    ms_e2e, win_b = time_e2e(lanes, args.steps, args.warmup, grid_mode, args.e2e_threads)
    ctx.set_option("async_counts", 1)
    e2e_check = float(sum(np.abs(o.array).sum() for _, o in lanes))   # the read-backs are real
    assert np.isfinite(e2e_check) and e2e_check > 0
    # for comparison, the same loop with the other layout of the position rows (N x 3 through pgt_frame_upload: 16 bytes per
    # particle up; N x 2 through pgt_frame_upload_xy: 12).  At one GPU the link is not the bound and the two agree; with 4-8
    # ranks sharing one host the bytes count
    e2e_other, e2e_alt = None, None
    if not grid_mode and world == 1:
        main_lanes = [(c, o) for c, o in lanes[:args.e2e_lanes]] if args.e2e_lanes > 1 else [lanes[0]]
        if args.e2e_lanes > 1:
            ctx.set_option("async_counts", 0)
        pos_xy[0] = not pos_xy[0]
        ms_o, _ = time_e2e(list(main_lanes), args.steps, args.warmup, False, args.e2e_threads)
        pos_xy[0] = not pos_xy[0]
        ctx.set_option("async_counts", 1)
        e2e_other = ms_o / args.steps
        # for comparison, the other arrangement: ONE context and one host thread pipelining upload / stages / read-back on
        # the context's three streams with deferred counts (when the default is several contexts), or three contexts with a
        # host thread each (when one context was asked for)
        if args.e2e_lanes > 1:
            ms_alt, _ = time_e2e([lanes[0]], args.steps, args.warmup, False, 1)
        else:
            alt = make_lanes(ctx, 3, my_shard)
            ctx.set_option("async_counts", 0)
            ms_alt, _ = time_e2e(alt, args.steps, args.warmup, False, 2)
            ctx.set_option("async_counts", 1)
            for c, _ in alt[1:]:
                c.close()
        e2e_alt = ms_alt / args.steps

    # ---- N > 1: one frame split over the ranks by grid points (the reference's MPI decomposition), strong scaling
    grid_shard = None
    if world > 1 and not grid_mode and not args.no_grid_shard:
        gsteps = max(10, args.steps // 2)
        for c in {id(c): c for c, _ in lanes if c is not ctx}.values():
            c.close()
        # every rank works on rank 0's frames now
        frames0 = make_frames(wl, 0, pool)
        for k, f in enumerate(frames0):
            pin_pos[k].array[...] = f["position"]
            pin_diam[k].array[...] = f["diameter"]
            dev_pos[k].copy_(torch.from_numpy(f["position"]))
            dev_diam[k].copy_(torch.from_numpy(f["diameter"]))
        torch.cuda.synchronize()
        full_ref = None
        if rank == 0:                                         # the single-GPU field of frame 0, for the bitwise check
            step_resident(ctx, 0)
            full_ref = ctx.ik_download()
        ctx.set_shard(*shards[rank])
        ms_g, launches_g, stage_g, _ = time_resident(ctx, gsteps, args.warmup)
        glanes = make_lanes(ctx, 1, shards[rank])
        comm_init_all([c for c, _ in glanes])
        ms_ge, _ = time_e2e(glanes, gsteps, args.warmup, True, 1)
        # the NCCL legs alone: frame all-gather, field all-gather
        step_resident(ctx, 0)
        gathered = torch.empty((G, 4), dtype=torch.float64, device="cuda")
        counts = [4 * (e - b) for b, e in shards]
        ctx.sync()
        barrier()
        ctx.event_record(4)
        for _ in range(10):
            ctx.comm_allgather_field(_capi.FIELD_VIRIAL, counts, gathered.data_ptr())
        ctx.event_record(5)
        ctx.sync()
        nccl_field_ms = max_over_ranks(ctx.event_elapsed_ms(4, 5) / 10)
        ctx.upload_part(pin_pos[0].array[p_first:p_last], pin_diam[0].array[p_first:p_last], p_first, N, box)
        ctx.sync()
        barrier()
        ctx.event_record(6)
        for _ in range(10):
            ctx.comm_allgather_frame()
        ctx.event_record(7)
        ctx.sync()
        nccl_frame_ms = max_over_ranks(ctx.event_elapsed_ms(6, 7) / 10)
        check = None
        if rank == 0:
            check = "bitwise" if np.array_equal(gathered.cpu().numpy(), full_ref) else "MISMATCH"
        step_resident(ctx, 0)
        kept = ctx.counts()["n_live"]
        t1_ms = ms_resident / args.steps                     # this run's single-GPU step on the same workload
        grid_shard = {
            "what": "ONE frame per step, its %d grid points split over %d GPUs at np.array_split boundaries (irvingkirkwood/__init__.py:53-60); every rank builds slab-local cell and bond lists" % (G, world),
            "value": G * gsteps / (ms_g * 1e-3), "unit": UNIT, "ms_per_step": ms_g / gsteps, "scaling": "strong",
            "efficiency_vs_n1": (t1_ms / (ms_g / gsteps)) / world,
            "stage_ms": {k: float(np.mean(v)) for k, v in stage_g.items()},
            "particles_kept_rank0": kept, "particles_total": N,
            "e2e": {"value": G * gsteps / (ms_ge * 1e-3), "unit": UNIT, "ms_per_step": ms_ge / gsteps,
                    "h2d_bytes_per_step": (p_last - p_first) * 16, "d2h_bytes_per_step": (shards[rank][1] - shards[rank][0]) * ncomp_out * 8,
                    "efficiency_vs_n1_e2e": ((ms_e2e / args.steps) / (ms_ge / gsteps)) / world,
                    "api": "pgt_frame_upload_part (1/%d of the frame over PCIe) + pgt_comm_allgather_frame (ncclAllGather over NVLink) .. pgt_ik_download_%sasync of the rank's own shard; one context, one host thread" % (world, "sym_" if ncomp_out == 3 else "")},
            "collective": "ncclAllGather (frame parts, %.1f MB per rank -> %.1f MB) inside the timed e2e region; ncclBroadcast group (final field gather, %.0f MB) timed separately" % ((p_last - p_first) * 16 / 1e6, N * 16 / 1e6, G * 32 / 1e6),
            "nccl_ms": {"allgather_frame": nccl_frame_ms, "allgather_field": nccl_field_ms},
            "gather_check": check, "steps": gsteps,
        }
        launches_resident += launches_g
    if rank == 0:
        sampler.stop()

    # ---- report ------------------------------------------------------------------------------
    units_per_step = G if not grid_mode else g_local
    total_units = sum_over_ranks(units_per_step * args.steps)
    value = total_units / (ms_resident * 1e-3)
    e2e_value = total_units / (ms_e2e * 1e-3)
    launches_total = int(sum_over_ranks(launches_resident))
    ik_avg_ms = float(np.mean([m for _, m in ik_ms]))
    flops_avg = float(np.mean([flops_per_frame[k] for k, _ in ik_ms])) * (g_local / G if grid_mode else 1.0)
    achieved = flops_avg / (ik_avg_ms * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs")
    # DRAM bytes of one gather launch: not a counter of THIS run -- the committed `ncu --set full` capture of the same
    # workload and kernel (profiles/), named so that a stale figure is visible
    traffic, traffic_src, traffic_commit = None, None, None
    try:
        if args.workload == "cfg5" and not grid_mode:
            prof = os.path.join("profiles", "r2_ik_quad_full_cfg5.txt")
            for line in open(os.path.join(ROOT, prof)):
                if line.startswith("# kernel source commit:"):
                    traffic_commit = line.split(":", 1)[1].split()[0]
                if line.startswith("traffic"):
                    traffic, traffic_src = float(line.split()[-1]), prof
    except Exception:
        pass
    h2d = N * (2 if pos_xy[0] else 3) * 4 + N * 4 if not grid_mode else (p_last - p_first) * 16
    d2h = g_local * ncomp_out * 8
    sb = stage_bytes(N, ncells, float(np.mean(hb_per_frame)))
    cells_ms, bonds_ms = float(np.mean(stage_ms["cells"])), float(np.mean(stage_ms["bonds"]))

    def stage_line(nbytes, ms):
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"bytes": nbytes, "ms": ms, "gbs": gbs, "frac": (gbs / hbm) if hbm else None}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_resident / args.steps, "higher_is_better": True,
            "scaling": "weak" if not grid_mode else "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": config_dict(args, wl, world, pool),
            "clocks": sampler.summary([win_a, win_b]),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "host_binding": numa, "api": ("C-ABI %s .. %%s, pinned host buffers; %%s" % ("pgt_frame_upload_xy (N x 2 float32 rows: a 2-D frame has no z column)" if pos_xy[0] else "pgt_frame_upload")) % ("pgt_ik_download_sym_async (xx,xy,yy)" if args.e2e_field == "sym" else "pgt_ik_download_async", "one context (upload, compute and read-back streams), one host thread, counts deferred" if args.e2e_lanes == 1 else "frames spread over %d contexts, %s" % (args.e2e_lanes, "one host thread per context" if args.e2e_threads == 2 else "one host thread"))},
            ("e2e_xyz_upload" if pos_xy[0] else "e2e_xy_upload"): None if e2e_other is None else {
                "value": G / (e2e_other * 1e-3), "unit": UNIT, "ms_per_step": e2e_other, "h2d_bytes_per_step": N * (16 if pos_xy[0] else 12),
                "api": "the same loop through %s, for comparison" % ("pgt_frame_upload with N x 3 float32 position rows (the GSD layout)" if pos_xy[0] else "pgt_frame_upload_xy (N x 2 float32 position rows)")},
            ("e2e_one_context" if args.e2e_lanes > 1 else "e2e_three_contexts"): None if e2e_alt is None else {
                "value": G / (e2e_alt * 1e-3), "unit": UNIT, "ms_per_step": e2e_alt,
                "api": ("for comparison: ONE context and one host thread, upload / stages / read-back pipelined on the context's three streams, counts deferred (no host synchronisation between stages)" if args.e2e_lanes > 1 else "for comparison: three contexts, one host thread each, every stage waits for its count")},
            "e2e_link": "%.0f MB per frame cross PCIe (H2D %.0f + D2H %.0f); the link itself moves that in ~2.0-2.1 ms with both directions busy (tools/link_probe.py: 55 / 56 GB/s alone, ~84 GB/s together), less than the stages take; what the end-to-end loop loses against the resident one is the slowdown of the memory-bound stages while the read-back's DMA runs beside them (tools/pipe_probe.py: cells 0.22 -> 0.32 ms, bonds 0.53 -> 0.65 ms, gather +0.02 ms)" % ((h2d + d2h) / 1e6, h2d / 1e6, d2h / 1e6),
            "gpu_launches": launches_total,
            "stage_ms": {"cells": cells_ms, "bonds": bonds_ms, "ik": ik_avg_ms},
            "roofline": {"kernel": "k_ik_quad", "bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp64_peak,
                         "peak_source": "DFMA-loop probe run on this GPU in this process (MEASURED_PEAKS.json has no FP64 entry); "
                                        "nominal 148 SM x 64 FMA/clk x 2 x 1.965 GHz = 37.2",
                         "frac_of_nominal_37.2": achieved / 37.2,
                         "work": "sum_x U(x)*(15q+9) flops, U counted on device (pgt_ik_count_work), %.4g flops per launch"
                                 % flops_avg,
                         "items_per_unique_bond": 1.0,
                         "items_note": "the quad gather integrates every unique bond once (weight 1 or 1/2); the reference's half-bond formulation visits %.3f items per unique bond"
                                       % (float(np.mean(items_per_frame)) / max(1.0, float(np.mean(flops_per_frame)) / (15 * q + 9))),
                         "traffic": traffic, "traffic_from_profile": traffic_src, "traffic_profile_kernel_commit": traffic_commit,
                         "hbm_peak_gbs": hbm},
            "roofline_stages": {"bound": "hbm", "peak_gbs": hbm, "peak_source": "MEASURED_PEAKS.json hbm_gbs (driver-measured copy bandwidth)",
                                "bytes_model": "SURVEY.md 8(d): cells 60 B/particle + 8 B/cell; bonds 32 B/particle + 36 B/half-bond",
                                "cells": stage_line(sb["cells"], cells_ms), "bonds": stage_line(sb["bonds"], bonds_ms)},
        }
        if grid_shard is not None:
            out["grid_shard"] = grid_shard
        if world == 1 and not args.no_other_configs and args.workload == "cfg5":
            oc = {}
            for key, fn in (("cfg1", lambda: fixture_workload(_capi, synth, local_rank)),
                            ("cfg2", lambda: thermo_workload(_capi, synth, local_rank)),
                            ("cfg3", lambda: quick_workload(_capi, synth, local_rank, "cfg3")),
                            ("cfg4", lambda: quick_workload(_capi, synth, local_rank, "cfg4"))):
                try:
                    oc[key] = fn()
                except Exception as e:                       # a side measurement must not lose the main line
                    oc[key] = {"error": repr(e)}
            try:                                             # the Python surface, text output included (child process)
                res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "script_rate.py"), "1024"], capture_output=True,
                                     text=True, timeout=300)
                oc.update(json.loads(res.stdout.strip().splitlines()[-1]))
            except Exception as e:
                oc["script"] = {"error": repr(e)}
            out["other_configs"] = oc
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(wl, frames[0], args)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(wl, fr, args):
    """The oracle on the box's host cores, on a bounded strided sample of the same frame's grid: the timing build
    (the reference's flags) is the reported number, the parity build's rate is given beside it once."""
    stride = args.cpu_sample_stride or max(1, int(round(wl["nmax"] / 1024.0)))
    r = oracle_rate(wl, fr, stride, 10.0, 1)
    out = {"value": r["rate"], "unit": UNIT, "cores": r["cores"], "kind": "port",
           "sample": "%d of %d grid points (every %d-th in i and j) of one frame, %d passes of %.1f s; oracle frame build %.1f s excluded; %s = oracle/pgt_oracle.cpp built -O3 -march=native -fopenmp (CompilerFlagsSetup.cmake:14)"
                     % (r["points"], wl["nmax"] ** 2, r["stride"], r["passes"], r["dt"], r["build_s"], r["lib"])}
    # the parity build (-O2 -ffp-contract=off) in a child process (one oracle library per process), a shorter sample
    try:
        code = ("import json,sys;sys.path.insert(0,%r);import bench;from pyglasstools_b200 import synth;"
                "import os;os.environ.pop('PGT_ORACLE_FAST',None);"
                "wl=bench.WORKLOADS[%r];fr=synth.synth_frame(wl['n'],synth.SEEDS[wl['seed']]);"
                "from oracle import orc;import numpy as np,time;orc.set_num_threads(len(os.sched_getaffinity(0)));"
                "n,rc,prm=bench.pot_params(wl['pot']);P=orc.potential(n);cg=orc.cgfunc(*wl['cg']);S=orc.System(fr['box'],fr['position'],fr['diameter']);"
                "ax=synth.linspace_axis(float(fr['box'][0]),wl['nmax']);ii=np.arange(0,wl['nmax'],%d);g=np.zeros((len(ii)**2,3));g[:,0]=np.repeat(ax[ii],len(ii));g[:,1]=np.tile(ax[ii],len(ii));"
                "mr=S.max_rcut(P);S.ik(P,cg,g[:4096],max_rcut=mr);t=time.perf_counter();S.ik(P,cg,g,max_rcut=mr);print(json.dumps(len(g)/(time.perf_counter()-t)))"
                % (ROOT, args.workload, stride * 2))
        env = dict(os.environ)
        env.pop("PGT_ORACLE_FAST", None)
        res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=180, env=env)
        out["parity_build_value"] = float(json.loads(res.stdout.strip().splitlines()[-1]))
        out["parity_build"] = "liborc.so (-O2 -ffp-contract=off, the ISO-C reading used by the parity tests), every %d-th grid row and column" % (stride * 2)
    except Exception as e:
        out["parity_build_value"] = None
        out["parity_build"] = "not measured: %r" % (e,)
    return out


if __name__ == "__main__":
    main()
