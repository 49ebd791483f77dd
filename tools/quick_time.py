"""Quick device timing of the stages on a synthetic frame (developer tool, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyglasstools_b200 import _capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
R = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
cell = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
fr = synth.synth_frame(n, synth.SEEDS["cfg3"])
ctx = _capi.Context(0)
print("fp64 peak probe TFLOP/s:", ctx.fp64_peak())
rc = 1.25
ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
ctx.set_cg("octic", R, 5)
if cell > 0:
    ctx.set_cell_size(cell)
ctx.set_max_rcut(max(rc * float(fr["diameter"].max()), float(fr["diameter"].max())))
L = float(fr["box"][0])
axis = synth.linspace_axis(L, n)
ctx.set_grid_cartesian(axis, axis)
for rep in range(reps):
    t0 = time.time()
    ctx.upload(fr["position"], fr["diameter"], fr["box"])
    ctx.build_cells(); ctx.build_bonds(); ctx.ik_compute(1); ctx.sync()
    t1 = time.time()
    print("rep", rep, "wall ms %.3f" % ((t1 - t0) * 1e3), {s: round(ctx.stage_ms(s), 4) for s in ("upload", "cells", "bonds", "ik")}, ctx.counts())
G = n * n
ms = ctx.stage_ms("ik")
print("ik: %.3e gridpoint/s" % (G / (ms * 1e-3)))
f = ctx.ik_download()
print("field mean", f.mean(0))
t = ctx.thermo(1 | 8); print("thermo ms", ctx.stage_ms("thermo"), t["virial"], t["energy"])
