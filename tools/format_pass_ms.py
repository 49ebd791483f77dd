import os, sys, time
sys.path.insert(0, "/root/repo/pyglasstools_b200")
import numpy as np
from pyglasstools import _pgt_host as H
from pyglasstools.irvingkirkwood import _irvingkirkwood
n = 1024
ax = np.linspace(-n / 2, n / 2, n)
pts = np.zeros((n * n, 3)); pts[:, 0] = np.repeat(ax, n); pts[:, 1] = np.tile(ax, n)
vals = np.random.default_rng(1).normal(size=(n * n, 4)) * 100.0
f = _irvingkirkwood.VirialStressField2D("virialstress", "2-TENSOR", False, len(pts))
f.addGridpoints(pts); f.set_values(vals)
for nt in (1, 4, 8, 16, 0):
    best = min(H.format_field_frame_ms(pts, 2, [f, f, f], [0, 3, 2], nt) for _ in range(4))
    print("threads %s: formatting pass alone %.1f ms" % (nt or "all", best))
