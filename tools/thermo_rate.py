"""cfg2 (BASELINE configs[1]): thermo global virial stress + potential energy, 2D polydisperse12, synthetic N = 65536,
frames of a trajectory streamed from pinned host memory: frames/s through the C-ABI (upload + cells + thermo sums +
25-double read-back per frame), next to the CPU oracle's serial loop on a few frames.  Developer tool, not the bench."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pyglasstools_b200 import _capi, synth

nframes = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pool = [synth.synth_frame(256, synth.SEEDS["cfg2"] + f) for f in range(8)]
ctx = _capi.Context(0)
rc = 1.25
ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
dmax = max(float(fr["diameter"].max()) for fr in pool)
ctx.set_max_rcut(max(rc * dmax, dmax))
mask = _capi.OBS_VIRIAL | _capi.OBS_ENERGY
for rep in range(3):
    t0 = time.perf_counter()
    for f in range(nframes):
        fr = pool[f % len(pool)]
        ctx.upload(fr["position"], fr["diameter"], fr["box"])
        ctx.build_cells()
        out = ctx.thermo(mask)
    dt = time.perf_counter() - t0
    print("rep %d: %d frames in %.3f s -> %.1f frames/s (cells %.3f ms, thermo %.3f ms per frame on the device)" % (
        rep, nframes, dt, nframes / dt, ctx.stage_ms("cells"), ctx.stage_ms("thermo")))
print("last frame: virial", out["virial"], "energy", out["energy"])
try:
    from oracle import orc
    P = orc.potential("poly12")
    fr = pool[(nframes - 1) % len(pool)]
    t0 = time.perf_counter()
    S = orc.System(fr["box"], fr["position"], fr["diameter"])
    w = S.thermo(P, born=False)
    dt = time.perf_counter() - t0
    print("oracle (serial CPU loop, ThermoCalculator.h:68-90): %.3f s per frame -> %.2f frames/s; |dvirial| %.2e" % (
        dt, 1 / dt, np.abs(w["virial"] - out["virial"]).max()))
except Exception as e:   # the oracle is test infrastructure: the tool still reports the device rate without it
    print("oracle unavailable:", e)
