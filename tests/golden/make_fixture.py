#!/usr/bin/env python
"""Extract the reference's only data fixture into a small array file (run in the build container).

Reads /root/reference/example_scripts/polydisperse12.gsd (201 frames, N=1024, 2D, float32) with the
GSD reader vendored by the reference (pyglasstools/extern/gsd/gsd/pygsd.py -- third-party, imported
from where it lies) and stores positions (x,y; z is identically 0), the frame-independent diameters
and the box as tests/golden/polydisperse12_frames.npz.  /root/reference does not exist on the GPU
box, so tests rebuild a GSD file from this archive with the repo's own writer.
"""
import importlib.util
import os

import numpy as np

SRC = "/root/reference/example_scripts/polydisperse12.gsd"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    spec = importlib.util.spec_from_file_location("pygsd", "/root/reference/pyglasstools/extern/gsd/gsd/pygsd.py")
    pygsd = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(pygsd)
    f = pygsd.GSDFile(open(SRC, "rb"))
    nf = f.nframes
    d0 = f.read_chunk(0, "particles/diameter")
    b0 = f.read_chunk(0, "configuration/box")
    xy = np.empty((nf, len(d0), 2), np.float32)
    for fr in range(nf):
        p = f.read_chunk(fr, "particles/position")
        assert p.dtype == np.float32 and np.all(p[:, 2] == 0)
        assert np.array_equal(f.read_chunk(fr, "particles/diameter"), d0)
        assert np.array_equal(f.read_chunk(fr, "configuration/box"), b0)
        assert not f.chunk_exists(fr, "particles/velocity") and not f.chunk_exists(fr, "particles/mass")
        xy[fr] = p[:, :2]
    np.savez_compressed(os.path.join(HERE, "polydisperse12_frames.npz"), xy=xy, diameter=d0, box=b0)
    print("frames", nf, "N", len(d0), "box", b0)


if __name__ == "__main__":
    main()
