import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np
from pyglasstools_b200 import _capi
ctx = _capi.Context(0)
for mb in (8, 32, 64):
    a = np.ones(mb * 131072)      # float64: mb MB, touched
    for rep in range(3):
        t0 = time.perf_counter(); ok = _capi.host_register(a); t1 = time.perf_counter(); _capi.host_unregister(a); t2 = time.perf_counter()
        print("%d MB: register %.2f ms (%s), unregister %.2f ms" % (mb, (t1 - t0) * 1e3, ok, (t2 - t1) * 1e3))
