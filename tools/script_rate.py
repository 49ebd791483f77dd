"""Script-level throughput: the reference's example script (example_scripts/irvingkirkwood/test-irvingkirkwood-module.py:11-25)
statement for statement through the drop-in Python surface -- read_gsd, ikcalculator(dx), fieldlogger (text .xyz output), analyze --
on (a) the shipped fixture trajectory, frames 180..199, and (b) a synthetic N = n^2 trajectory written to GSD with a
(n x n) grid.  Prints one JSON object.  Used by bench.py (other_configs) in a child process; needs a GPU."""
import json, os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "pyglasstools_b200"))
import numpy as np


def run_script(gsd_path, dx, frames):
    import pyglasstools as pglass
    import pyglasstools.io as io
    import pyglasstools.utils as utils
    import pyglasstools.potential as pot
    import pyglasstools.irvingkirkwood as ik
    import pyglasstools.irvingkirkwood.cgfunc as cgfunc
    pglass.reset()
    poly12 = pot.polydisperse12(v0=1.0, eps=0.2, rcut=1.25)          # noqa: F841
    mycgfunc = cgfunc.octic(rcut=2.0, mode="fixedpoint", order=5)
    mysystem = utils.read_gsd(filename=gsd_path, checkpointfile="restart.log")   # noqa: F841
    calc = ik.ikcalculator(cgfunc=mycgfunc, dx=dx)
    logger = io.fieldlogger(keyword="dump", names=['virialstress_00', 'virialstress_11', 'virialstress_01'], solver=calc)
    # where the time of a frame goes: frame read + particle build / device compute + read-back into the holders / text output
    split = {"update": 0.0, "run": 0.0, "save": 0.0}

    def timed(obj, name, key):
        f = getattr(obj, name)

        def g(*a, **k):
            t = time.perf_counter()
            r = f(*a, **k)
            split[key] += time.perf_counter() - t
            return r
        setattr(obj, name, g)
    timed(calc, "update", "update")
    timed(calc, "run", "run")
    timed(logger, "save", "save")
    t0 = time.perf_counter()
    pglass.analyze(frame_list=frames)
    logger.file.flush()
    dt = time.perf_counter() - t0
    run_script.split = {k: v / max(1, len(frames)) * 1e3 for k, v in split.items()}
    G = calc.gridsize
    size = os.path.getsize("dump.xyz")
    pglass.reset()
    return G, dt, size


def main():
    from pyglasstools import gsdio
    from pyglasstools_b200 import synth
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    out = {}
    with tempfile.TemporaryDirectory() as d:
        os.chdir(d)
        z = np.load(os.path.join(ROOT, "tests", "golden", "polydisperse12_frames.npz"))
        frames = []
        for f in range(z["xy"].shape[0]):
            p = np.zeros((z["xy"].shape[1], 3), np.float32)
            p[:, :2] = z["xy"][f]
            frames.append(dict(box=z["box"], position=p, diameter=z["diameter"]))
        gsdio.write_trajectory("fixture.gsd", frames)
        run_script("fixture.gsd", 0.25, range(180, 183))            # warm-up (context, allocations)
        os.remove("restart.log")
        G, dt, size = run_script("fixture.gsd", 0.25, range(180, 200))
        out["cfg1_script"] = {"description": "example_scripts/irvingkirkwood on the shipped trajectory, frames 180..199, dx = 0.25 (grid %d), text .xyz output included" % G,
                              "value": G * 20 / dt, "unit": "grid-point*frames/s", "ms_per_frame": dt / 20 * 1e3, "text_MB_per_frame": size / 20 / 1e6, "ms_split": run_script.split}
        os.remove("restart.log"); os.remove("dump.xyz")
        nfr = 10
        syn = [synth.synth_frame(n, 777 + k) for k in range(nfr)]
        gsdio.write_trajectory("synth.gsd", [dict(box=f["box"], position=f["position"], diameter=f["diameter"]) for f in syn])
        L = float(syn[0]["box"][0])
        run_script("synth.gsd", L / n, range(0, 1))
        os.remove("restart.log"); os.remove("dump.xyz")
        G, dt, size = run_script("synth.gsd", L / n, range(0, nfr))
        out["synthetic_script"] = {"description": "the same script on a synthetic N = %d trajectory written to GSD, grid %d, %d frames, text .xyz output included" % (n * n, G, nfr),
                                   "value": G * nfr / dt, "unit": "grid-point*frames/s", "ms_per_frame": dt / nfr * 1e3, "text_MB_per_frame": size / nfr / 1e6, "ms_split": run_script.split}
        # BASELINE configs[1] at script level: example_scripts/thermo/test-thermo-module.py's statements on a synthetic
        # N = 65536 trajectory (virial stress + potential energy into a log file)
        os.remove("restart.log"); os.remove("dump.xyz")
        nth = 200
        pool = [synth.synth_frame(256, 4000 + k) for k in range(8)]
        gsdio.write_trajectory("thermo.gsd", [dict(box=pool[k % 8]["box"], position=pool[k % 8]["position"], diameter=pool[k % 8]["diameter"])
                                              for k in range(nth)])
        import pyglasstools as pglass
        import pyglasstools.io as io
        import pyglasstools.utils as utils
        import pyglasstools.potential as pot
        import pyglasstools.thermo as thermo
        for frames_ in (range(0, 8), range(0, nth)):
            pglass.reset()
            if os.path.exists("restart.log"):
                os.remove("restart.log")
            poly12 = pot.polydisperse12(v0=1.0, eps=0.2, rcut=1.25)          # noqa: F841
            mysystem = utils.read_gsd(filename="thermo.gsd", checkpointfile="restart.log")   # noqa: F841
            calc = thermo.calculator()
            log = io.logfile(filename="thermo.log", names=['virialstress_00', 'virialstress_11', 'virialstress_01', 'potentialenergy'], solver=calc)
            split = {"update": 0.0, "run": 0.0, "save": 0.0}

            def timed(obj, name, key):
                f = getattr(obj, name)

                def g(*a, **k):
                    t = time.perf_counter()
                    r = f(*a, **k)
                    split[key] += time.perf_counter() - t
                    return r
                setattr(obj, name, g)
            timed(calc, "update", "update")
            timed(calc, "run", "run")
            timed(log, "save", "save")
            t0 = time.perf_counter()
            pglass.analyze(frame_list=frames_)
            dt = time.perf_counter() - t0
            del log
            pglass.reset()
        out["cfg2_script"] = {"description": "example_scripts/thermo on a synthetic N = 65536 trajectory written to GSD, %d frames, virial stress + potential energy into the log file" % nth,
                              "value": nth / dt, "unit": "frames/s", "ms_per_frame": dt / nth * 1e3,
                              "ms_split": {k: v / nth * 1e3 for k, v in split.items()}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
