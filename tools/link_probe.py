"""Host-link probe (developer tool): D2H / H2D rates of pinned buffers alone, together, and while the stages run."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pyglasstools_b200 import _capi, synth

n = 2048
dev = torch.device("cuda:0")
d_out = torch.empty(n * n * 3, dtype=torch.float64, device=dev)
h_out = [torch.empty(n * n * 3, dtype=torch.float64).pin_memory() for _ in range(3)]
h_in = torch.empty(n * n * 4, dtype=torch.float32).pin_memory()
d_in = torch.empty(n * n * 4, dtype=torch.float32, device=dev)
s_dn, s_up = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=20):
    fn(); torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.time() - t0) / reps * 1e3


def dn():
    with torch.cuda.stream(s_dn):
        h_out[0].copy_(d_out, non_blocking=True)


def up():
    with torch.cuda.stream(s_up):
        d_in.copy_(h_in, non_blocking=True)


def both():
    dn(); up()


def report(tag):
    a, b, c = timed(dn), timed(up), timed(both)
    print("%-28s D2H 100.7 MB %.3f ms (%.1f GB/s) | H2D 67.1 MB %.3f ms (%.1f GB/s) | both %.3f ms (%.1f GB/s total)" % (
        tag, a, 100.7 / a, b, 67.1 / b, c, 167.8 / c))


report("idle GPU")
fr = synth.synth_frame(n, 1)
ctx = _capi.Context(0)
rc = 1.25
ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
ctx.set_cg("octic", 2.0, 5)
ctx.set_max_rcut(max(rc * float(fr["diameter"].max()), float(fr["diameter"].max())))
axis = synth.linspace_axis(float(fr["box"][0]), n)
ctx.set_grid_cartesian(axis, axis)
dp = torch.from_numpy(fr["position"]).cuda(); dd = torch.from_numpy(fr["diameter"]).cuda()
stop = [False]


def work():
    while not stop[0]:
        for _ in range(8):
            ctx.set_device_frame(dp.data_ptr(), dd.data_ptr(), n * n, fr["box"], is_f32=True)
            ctx.build_cells(); ctx.build_bonds(); ctx.ik_compute(1)
        ctx.sync()


t = threading.Thread(target=work); t.start()
time.sleep(0.5)
report("stages running")
stop[0] = True; t.join()
