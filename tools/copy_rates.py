"""Measured host<->device copy rates of the C-ABI's pinned buffers (developer tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyglasstools_b200 import _capi, synth

n = 2048
fr = synth.synth_frame(n, 1)
ctx = _capi.Context(0)
rc = 1.25
ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
ctx.set_cg("octic", 2.0, 5)
ctx.set_max_rcut(max(rc * float(fr["diameter"].max()), float(fr["diameter"].max())))
axis = synth.linspace_axis(float(fr["box"][0]), n)
ctx.set_grid_cartesian(axis, axis)
pp = _capi.PinnedArray((n * n, 3), np.float32); pd = _capi.PinnedArray((n * n,), np.float32)
pp.array[...] = fr["position"]; pd.array[...] = fr["diameter"]
po = _capi.PinnedArray((n * n, 4), np.float64)
for rep in range(4):
    ctx.upload(pp.array, pd.array, fr["box"]); ctx.build_cells(); ctx.build_bonds(); ctx.ik_compute(1)
    ctx.ik_download(0, out=po.array)
    up, dn = ctx.stage_ms("upload"), ctx.stage_ms("download")
    print("H2D %.1f MB in %.3f ms = %.1f GB/s ; D2H %.1f MB in %.3f ms = %.1f GB/s" % (
        n * n * 16 / 1e6, up, n * n * 16 / up / 1e6, n * n * 32 / 1e6, dn, n * n * 32 / dn / 1e6))
