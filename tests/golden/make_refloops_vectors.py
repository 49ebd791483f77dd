#!/usr/bin/env python
"""Golden vectors from the REFERENCE'S OWN calculators (run in the build container, where /root/reference exists).

oracle/_ref/libpgt_refloops.so is IrvingKirkwood::compute (irvingkirkwood/IrvingKirkwood.h:74-110),
ThermoCalculator::compute (thermo/ThermoCalculator.h:68-90), ListofCommonObservables.h, CoarseGrainFunction.h,
FixedPointQuadrature.h, CoarseGrainedField.h, ThermoProperty.h, ParticleSystem.cc and SimBox.cc compiled UNMODIFIED from
where they lie, against the stand-ins of oracle/stubs/ (Aboria search, GSL rules, MPI log file) -- oracle/Makefile,
target `ref`.  This script runs them on the shipped fixture frames and on a small seam-crossing system and stores inputs
and outputs as tests/golden/refloops_vectors.npz, so that the GPU box (no /root/reference) can pin the oracle's loop
nests -- and through them the CUDA path -- to the reference's code:

  ik_<case>_{grid,virial}     IrvingKirkwood::compute with VirialStressField2D on the case's grid points
  ik_<case>_meta              frame index (or -1), kernel id, cg_rcut, order, max_rcut the calculator froze
  th_<frame>_{virial,energy,born,max_rcut}   ThermoCalculator::compute (already times 1/vol)
  small_{box,pos,diam}        the 64-particle seam-crossing system
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import orc  # noqa: E402

KERNEL_ID = {"octic": 0, "rect": 1, "mollifier": 2}
# (case name, frame or "small", potential, kernel, cg_rcut, order, grid points per axis)
IK_CASES = [
    ("f180_octic5", 180, "poly12", "octic", 2.0, 5, 24),
    ("f000_octic5", 0, "poly12", "octic", 2.0, 5, 16),
    ("f199_octic3", 199, "poly12", "octic", 1.5, 3, 16),
    ("f180_rect5", 180, "poly12", "rect", 2.0, 5, 16),
    ("f180_moll4", 180, "poly12", "mollifier", 2.5, 4, 16),
    ("f180_poly10_R4", 180, "poly10", "octic", 4.0, 5, 12),
    ("small_octic5", "small", "poly12", "octic", 2.0, 5, 13),
    ("small_lj_rect2", "small", "lj", "rect", 1.0, 2, 13),
]
THERMO_FRAMES = [0, 90, 180, 199]


def small_system():
    """64 particles in an 8 x 8 box, some exactly on the box faces, some outside (wrapped by init_neighbour_search)."""
    rng = np.random.default_rng(11)
    L = 8.0
    pos = np.zeros((64, 3))
    pos[:, :2] = rng.uniform(-L / 2, L / 2, (64, 2))
    pos[0, :2] = (-L / 2, 0.25)              # on the lower face
    pos[1, :2] = (L / 2, -0.5)               # on the upper face: wraps to the lower one
    pos[2, :2] = (L / 2 + 0.75, L / 2 + 0.25)  # outside on both axes
    pos[3, :2] = (-L / 2 - 2.0 * L, 1.0)     # two box lengths outside
    diam = rng.uniform(0.8, 1.3, 64)
    return np.array([L, L, 1.0]), pos, diam


def grid_points(L, n):
    ax = np.linspace(-L / 2, L / 2, n)       # both ends included, like the ikcalculator lattice
    g = np.zeros((n * n, 3))
    g[:, 0] = np.repeat(ax, n)
    g[:, 1] = np.tile(ax, n)
    return g


def main():
    if orc.refloops_lib() is None:
        raise SystemExit("oracle/_ref/libpgt_refloops.so is not built (needs /root/reference): make -C oracle ref")
    z = np.load(os.path.join(HERE, "polydisperse12_frames.npz"))
    out = {}
    sbox, spos, sdiam = small_system()
    out["small_box"], out["small_pos"], out["small_diam"] = sbox, spos, sdiam
    for name, frame, pot, kernel, R, order, n in IK_CASES:
        if frame == "small":
            box, pos, diam = sbox, spos, sdiam
            fidx = -1
        else:
            pos = np.zeros((z["xy"].shape[1], 3), np.float32)
            pos[:, :2] = z["xy"][frame]
            box, diam, fidx = z["box"], z["diameter"], frame
        P = orc.potential(pot)
        grid = grid_points(float(box[0]), n)
        r = orc.ref_ik(box, pos, diam, P, kernel, R, order, grid)
        out["ik_%s_grid" % name] = grid
        out["ik_%s_virial" % name] = r["virial"]
        out["ik_%s_meta" % name] = np.array([fidx, KERNEL_ID[kernel], R, order, r["max_rcut"], r["n_live"]], np.float64)
    P = orc.potential("poly12")
    for f in THERMO_FRAMES:
        pos = np.zeros((z["xy"].shape[1], 3), np.float32)
        pos[:, :2] = z["xy"][f]
        t = orc.ref_thermo(z["box"], pos, z["diameter"], P)
        out["th_%03d_virial" % f] = t["virial"]
        out["th_%03d_energy" % f] = np.array([t["energy"]])
        out["th_%03d_born" % f] = t["born"]
        out["th_%03d_max_rcut" % f] = np.array([t["max_rcut"]])
    path = os.path.join(HERE, "refloops_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%d arrays, %.1f kB)" % (path, len(out), os.path.getsize(path) / 1e3))


if __name__ == "__main__":
    main()
