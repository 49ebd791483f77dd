"""Frames of a resident trajectory spread over several contexts (= compute streams) of one GPU: the memory-bound stages of one
frame can run beside the gather of another (developer tool)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pyglasstools_b200 import _capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
pool = 3
frames = [synth.synth_frame(n, 10 + k) for k in range(pool)]
rc = 1.25
axis = synth.linspace_axis(float(frames[0]["box"][0]), n)
box = frames[0]["box"]
dmax = max(float(f["diameter"].max()) for f in frames)


def mk():
    c = _capi.Context(0)
    c.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    c.set_cg("octic", 2.0, 5)
    c.set_max_rcut(max(rc * dmax, dmax))
    c.set_grid_cartesian(axis, axis)
    return c


dpos = [torch.from_numpy(f["position"]).cuda() for f in frames]
ddiam = [torch.from_numpy(f["diameter"]).cuda() for f in frames]
ctxs = [mk() for _ in range(4)]


def step(c, k):
    c.set_device_frame(dpos[k].data_ptr(), ddiam[k].data_ptr(), n * n, box, is_f32=True)
    c.build_cells(); c.build_bonds(); c.ik_compute(1)


for nctx in (1, 2, 3, 4):
    use = ctxs[:nctx]
    for s in range(4 * nctx):
        step(use[s % nctx], s % pool)
    for c in use:
        c.sync()
    steps = 120
    t0 = time.perf_counter()
    for s in range(steps):
        step(use[s % nctx], s % pool)
    for c in use:
        c.sync()
    ms = (time.perf_counter() - t0) / steps * 1e3
    print("contexts %d : %.3f ms/frame  (%.4e grid-point*frames/s)" % (nctx, ms, n * n / ms * 1e3))
