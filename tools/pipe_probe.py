"""Where does the single-context end-to-end frame time go?  The bench's pipelined loop with the copies switched off one at
a time (developer tool)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pyglasstools_b200 import _capi, synth

n = 2048
pool = 3
frames = [synth.synth_frame(n, 10 + k) for k in range(pool)]
rc = 1.25
axis = synth.linspace_axis(float(frames[0]["box"][0]), n)
box = frames[0]["box"]
c = _capi.Context(0)
c.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
c.set_cg("octic", 2.0, 5)
dmax = max(float(f["diameter"].max()) for f in frames)
c.set_max_rcut(max(rc * dmax, dmax))
c.set_grid_cartesian(axis, axis)
pxy = [_capi.PinnedArray((n * n, 2), np.float32) for _ in range(pool)]
pd = [_capi.PinnedArray((n * n,), np.float32) for _ in range(pool)]
dpos, ddiam = [], []
for k, f in enumerate(frames):
    pxy[k].array[...] = f["position"][:, :2]; pd[k].array[...] = f["diameter"]
    dpos.append(torch.from_numpy(f["position"]).cuda()); ddiam.append(torch.from_numpy(f["diameter"]).cuda())
outs = [_capi.PinnedArray((n * n, 3), np.float64) for _ in range(4)]


def run(steps, h2d, d2h, nbuf, order):
    def up(s):
        k = s % pool
        if h2d:
            c.upload_xy(pxy[k].array, pd[k].array, box)
        else:
            c.set_device_frame(dpos[k].data_ptr(), ddiam[k].data_ptr(), n * n, box, is_f32=True)
    up(0)
    for s in range(steps):
        c.build_cells(); c.build_bonds()
        if order == 1 and s + 1 < steps:
            up(s + 1)
        c.ik_compute(1)
        if order == 0 and s + 1 < steps:
            up(s + 1)
        if d2h:
            c.ik_download_sym_async(0, outs[s % nbuf].array)
            c.ik_download_wait_keep(nbuf - 1)
    c.ik_download_wait()


for order in (0,):
    for h2d, d2h, nbuf in ((1, 1, 3), (1, 1, 2), (1, 1, 4), (1, 0, 3), (0, 1, 3), (0, 0, 3)):
        run(6, h2d, d2h, nbuf, 0); c.sync()
        c.event_record(0)
        t0 = time.perf_counter()
        run(40, h2d, d2h, nbuf, 0)
        c.event_record(1); c.sync()
        print("order %d h2d %d d2h %d buffers %d : %.3f ms/frame (device) %.3f (host)  last frame's stages %s" % (
            order, h2d, d2h, nbuf, c.event_elapsed_ms(0, 1) / 40, (time.perf_counter() - t0) / 40 * 1e3,
            {k: round(c.stage_ms(k), 3) for k in ("cells", "bonds", "ik")}))
