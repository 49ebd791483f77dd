"""Throughput of the field text formatter (io.fieldlogger.save_perrank) on this host: serial and all threads, next to the
round-1 cost model (one snprintf("%g") per value through ostream_double).  Developer tool; no GPU needed."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pyglasstools_b200"))
import numpy as np
from pyglasstools import _pyglasstools
from pyglasstools.irvingkirkwood import _irvingkirkwood

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ax = np.linspace(-n / 2, n / 2, n)
pts = np.zeros((n * n, 3)); pts[:, 0] = np.repeat(ax, n); pts[:, 1] = np.tile(ax, n)
vals = np.random.default_rng(1).normal(size=(n * n, 4)) * 100.0
f = _irvingkirkwood.VirialStressField2D("virialstress", "2-TENSOR", False, len(pts))
f.addGridpoints(pts); f.set_values(vals)
for nt in (1, 2, 4, 8, 0):
    best = 1e9
    for rep in range(3):
        t = time.perf_counter(); b = _pyglasstools.format_field_frame_bytes(pts, 2, [f, f, f], [0, 3, 2], nt); best = min(best, time.perf_counter() - t)
    print("G = %d, %s threads: %.1f ms per frame (%.0f MB text)" % (n * n, nt or "all (%d)" % os.cpu_count(), best * 1e3, len(b) / 1e6))
t = time.perf_counter()
m = 200000
for v in vals[:m, 0]:
    _pyglasstools.ostream_double(float(v))
print("reference route, one Python -> C++ call per value (gettostring): %.0f ns per value -> %.1f s per frame of %d x 3 values"
      % ((time.perf_counter() - t) / m * 1e9, (time.perf_counter() - t) / m * n * n * 3, n * n))
