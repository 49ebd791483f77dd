"""Where does the end-to-end frame time go?  Variants of the double-buffered loop (developer tool)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyglasstools_b200 import _capi, synth

n = 2048
fr = synth.synth_frame(n, 1)
rc = 1.25
axis = synth.linspace_axis(float(fr["box"][0]), n)
def mk():
    ctx = _capi.Context(0)
    ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    ctx.set_cg("octic", 2.0, 5)
    ctx.set_max_rcut(max(rc * float(fr["diameter"].max()), float(fr["diameter"].max())))
    ctx.set_grid_cartesian(axis, axis)
    return ctx
pp = _capi.PinnedArray((n * n, 3), np.float32); pd = _capi.PinnedArray((n * n,), np.float32)
pp.array[...] = fr["position"]; pd.array[...] = fr["diameter"]
lanes = [(mk(), _capi.PinnedArray((n * n, 4), np.float64)) for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2)]

def run(nl, h2d, d2h, steps=20):
    def step(s):
        c, out = lanes[s % nl]
        if h2d or s < nl:
            c.upload(pp.array, pd.array, fr["box"])
        else:
            c.set_device_frame_same()
        c.build_cells(); c.build_bonds(); c.ik_compute(1)
        if d2h:
            c.ik_download_wait(); c.ik_download_async(0, out.array)
    for s in range(4): step(s)
    for c, _ in lanes: c.sync()
    t0 = time.perf_counter()
    for s in range(steps): step(s)
    for c, _ in lanes: c.sync()
    return (time.perf_counter() - t0) / steps * 1e3

# a frame that is already resident: re-run the stages without a new upload
def _same(self):
    lib = _capi.lib()
_capi.Context.set_device_frame_same = lambda self: None
for nl in range(1, len(lanes) + 1):
    for h2d, d2h in ((1, 1), (1, 0), (0, 1), (0, 0)):
        print("lanes %d  h2d %d  d2h %d : %.3f ms/frame" % (nl, h2d, d2h, run(nl, h2d, d2h)))
