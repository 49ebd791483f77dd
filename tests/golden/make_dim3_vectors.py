#!/usr/bin/env python
"""Golden vectors for the reference's 3-D typedefs (module-irvingkirkwood.cpp:9-13, module-thermo.cpp:8-20), produced by
the REFERENCE'S OWN calculators (oracle/_ref/libpgt_refloops.so: IrvingKirkwood::compute and ThermoCalculator::compute
compiled unmodified with Dim = 3, see tests/golden/make_refloops_vectors.py for what the stand-ins replace).  Run in the
build container, where /root/reference exists; writes tests/golden/dim3_vectors.npz:

  <case>_{box,pos,diam,vel,mass,disp,grid}            inputs (positions partly outside the box: wrapped on all three axes)
  <case>_{virial,elastic}                             G x 9 fields of VirialStressField3D / ElasticVirialStressField3D
  <case>_th                                           virial 9 | kinetic 9 | energy | born 81 | elastic 9, times 1/vol
  <case>_meta                                         potential id, kernel id, cg_rcut, order, max_rcut frozen by the calculator
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import orc  # noqa: E402

KERNEL_ID = {"octic": 0, "rect": 1, "mollifier": 2}
# name, particles per axis (x, y, z), box, potential, kernel, cg_rcut, order, grid points per axis, z of the grid plane
CASES = [
    ("cube_poly12_octic5", (6, 6, 6), (6.0, 6.0, 6.0), "poly12", "octic", 1.5, 5, 7, 0.0),
    ("slab_lj_rect3", (7, 5, 4), (7.7, 5.5, 4.4), "lj", "rect", 1.25, 3, 6, 0.0),
    ("brick_poly10_moll4", (5, 6, 7), (5.5, 6.6, 7.7), "poly10", "mollifier", 1.75, 4, 6, 0.4),
]


def system(nxyz, box, seed):
    rng = np.random.default_rng(seed)
    nx, ny, nz = nxyz
    L = np.asarray(box, np.float64)
    ii, jj, kk = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    frac = np.stack([(ii.ravel() + 0.5) / nx, (jj.ravel() + 0.5) / ny, (kk.ravel() + 0.5) / nz], 1)
    pos = (frac - 0.5) * L + rng.uniform(-0.3, 0.3, frac.shape) * (L / np.array([nx, ny, nz]))
    pos = pos.astype(np.float32).astype(np.float64)
    pos[1] += (L[0], 0.0, 0.0)                     # outside: wrapped on x
    pos[2] += (0.0, -2.0 * L[1], L[2])             # two boxes out on y, one on z
    pos[3, 2] = L[2] / 2                           # on the upper z face: wraps to the lower one
    n = len(pos)
    diam = rng.uniform(0.8, 1.3, n).astype(np.float32).astype(np.float64)
    vel = rng.standard_normal((n, 3))
    mass = 1.0 + rng.integers(0, 3, n).astype(np.float64)
    disp = 0.05 * rng.standard_normal((n, 3))
    return L, pos, diam, vel, mass, disp


def plane_grid(box, n, z):
    ax = np.linspace(-box[0] / 2, box[0] / 2, n)   # the ikcalculator lattice (irvingkirkwood/__init__.py:45-52): both axes from L_x
    g = np.zeros((n * n, 3))
    g[:, 0] = np.repeat(ax, n)
    g[:, 1] = np.tile(ax, n)
    g[:, 2] = z
    return g


def main():
    if orc.refloops_lib() is None:
        raise SystemExit("oracle/_ref/libpgt_refloops.so is not built (needs /root/reference)")
    out = {}
    for seed, (name, nxyz, box, pot, kernel, R, order, ng, z) in enumerate(CASES):
        L, pos, diam, vel, mass, disp = system(nxyz, box, 100 + seed)
        P = orc.potential(pot)
        grid = plane_grid(L, ng, z)
        ik = orc.ref_ik(L, pos, diam, P, kernel, R, order, grid, vel=vel, mass=mass, obs_mask=1 | 8, dim=3, disp=disp)
        th = orc.ref_thermo(L, pos, diam, P, vel=vel, mass=mass, obs_mask=1 | 2 | 4 | 8 | 16, dim=3, disp=disp)
        for k, v in (("box", L), ("pos", pos), ("diam", diam), ("vel", vel), ("mass", mass), ("disp", disp), ("grid", grid),
                     ("virial", ik["virial"]), ("elastic", ik["elastic"])):
            out["%s_%s" % (name, k)] = v
        out[name + "_th"] = np.concatenate([th["virial"], th["kinetic"], [th["energy"]], th["born"], th["elastic"]])
        out[name + "_meta"] = np.array([P.model, KERNEL_ID[kernel], R, order, ik["max_rcut"], ik["n_live"]], np.float64)
        print(name, "n_live", ik["n_live"], "max |virial|", np.abs(ik["virial"]).max(), "energy/vol", th["energy"])
    np.savez_compressed(os.path.join(HERE, "dim3_vectors.npz"), **out)


if __name__ == "__main__":
    main()
