import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = os.path.join(ROOT, "pyglasstools_b200")      # makes `import pyglasstools` resolve to the drop-in package
if PKG not in sys.path:
    sys.path.insert(0, PKG)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def fixture_frames():
    """The reference's shipped trajectory (example_scripts/polydisperse12.gsd) as arrays."""
    z = np.load(os.path.join(GOLDEN, "polydisperse12_frames.npz"))

    class Frames:
        xy = z["xy"]
        diameter = z["diameter"]
        box = z["box"]
        nframes = z["xy"].shape[0]

        @staticmethod
        def position(f):
            p = np.zeros((z["xy"].shape[1], 3), np.float32)
            p[:, :2] = z["xy"][f]
            return p

    return Frames


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def gpu_ctx():
    from pyglasstools_b200 import _capi
    ctx = _capi.Context(0)          # raises without a CUDA device: there is no fallback
    yield ctx
    ctx.close()


@pytest.fixture(scope="session")
def fixture_gsd(tmp_path_factory, fixture_frames):
    """The shipped trajectory re-written as a GSD file by our own writer (the reference tree is absent on the GPU box)."""
    from pyglasstools import gsdio
    path = str(tmp_path_factory.mktemp("gsd") / "polydisperse12.gsd")
    frames = [dict(box=fixture_frames.box, position=fixture_frames.position(f), diameter=fixture_frames.diameter)
              for f in range(fixture_frames.nframes)]
    gsdio.write_trajectory(path, frames)
    return path
