#!/usr/bin/env python
"""Golden vectors from the REFERENCE'S OWN headers (run in the build container, where /root/reference exists).

oracle/_ref/libpgt_refmath.so is the reference's potential/*.h and irvingkirkwood/cgfunc/CGFunc*.h compiled
from where they lie (oracle/Makefile, target `ref`).  This script evaluates them on fixed tables and
stores the outputs as tests/golden/refmath_vectors.npz, so that the GPU box (no /root/reference) can pin
the oracle -- and through it the CUDA path -- to the reference's arithmetic bit for bit:

  pair_<model>_{rij,di,dj}                 inputs (incl. separations straddling the pair cut-off)
  pair_<model>_{energy,fdivr,stiff,rcutsq} getPairEnergy / getPairForceDivR / getBondStiffness / getRcut
                                           (potential/PairPotential.h:92-122 + the model headers)
  cg_<kernel>_R<rcut>_{d2,delta}           Octic/Rect/Mollifier::compute(d2) (cgfunc/CGFunc*.h)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import orc  # noqa: E402

MODELS = ["lj", "fslj", "poly12", "poly18", "poly10", "poly106", "polyyukawa", "polylj"]
CG = [("octic", 2.0), ("octic", 4.0), ("rect", 2.0), ("rect", 1.5), ("mollifier", 2.0), ("mollifier", 3.0)]


def pair_inputs(P, seed):
    rng = np.random.default_rng(seed)
    n = 512
    di = rng.uniform(0.73, 1.62, n)
    dj = rng.uniform(0.73, 1.62, n)
    eps = P.p[2] if P.model == orc.MODEL_IDS["polyyukawa"] else P.p[1]
    sigma = 0.5 * (di + dj) * (1 - eps * np.abs(di - dj))
    rc = P.scaled_rcut * sigma
    r = rng.uniform(0.45, 1.15, n) * rc                        # inside and a bit outside the cut-off
    # a block that straddles the cut-off by a few ulp on either side
    k = np.arange(64)
    r[:64] = rc[:64] * (1.0 + (k - 32) * 2.0 ** -52)
    th = rng.uniform(0, 2 * np.pi, n)
    rij = np.zeros((n, 3))
    rij[:, 0] = r * np.cos(th)
    rij[:, 1] = r * np.sin(th)
    # float32-promoted diameters like the fixture's (ParticleSystem.cc:199-204)
    di[128:256] = di[128:256].astype(np.float32)
    dj[128:256] = dj[128:256].astype(np.float32)
    return rij, di, dj


def main():
    assert orc.ref_lib() is not None, "oracle/_ref/libpgt_refmath.so is missing: run `make -C oracle ref` where /root/reference exists"
    out = {}
    for m, name in enumerate(MODELS):
        P = orc.potential(name)
        rij, di, dj = pair_inputs(P, 1000 + m)
        t = orc.ref_pair_table(P, rij, di, dj)
        out["pair_%s_rij" % name] = rij
        out["pair_%s_di" % name] = di
        out["pair_%s_dj" % name] = dj
        for k, v in t.items():
            out["pair_%s_%s" % (name, k)] = v
    for kernel, R in CG:
        d2 = np.concatenate([np.linspace(0.0, 1.2 * R * R, 301), R * R * (1.0 + (np.arange(33) - 16) * 2.0 ** -52)])
        out["cg_%s_R%g_d2" % (kernel, R)] = d2
        out["cg_%s_R%g_delta" % (kernel, R)] = orc.ref_cg_table(kernel, R, d2)
    path = os.path.join(HERE, "refmath_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "with", len(out), "arrays")


if __name__ == "__main__":
    main()
