"""The reference's 3-D typedefs (VirialStressField3D, ElasticVirialStressField3D, KineticStressField3D,
GlobalVirialStress3D, GlobalKineticStress3D, GlobalPotentialEnergy3D, GlobalBornTensor3D, GlobalElasticVirialStress3D;
module-irvingkirkwood.cpp:9-13, module-thermo.cpp:8-20).

CPU: the oracle with dim = 3 against golden vectors produced by the reference's own calculators compiled with Dim = 3
(tests/golden/make_dim3_vectors.py).  GPU: the device path (pgt_dim3.cu through the C-ABI) against the oracle on the same
inputs, BASELINE.json's tolerance per component.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

CASES = ["cube_poly12_octic5", "slab_lj_rect3", "brick_poly10_moll4"]
POT_NAME = {0: "lj", 2: "poly12", 4: "poly10"}
KERNEL = {0: "octic", 1: "rect", 2: "mollifier"}
GPU_POT = {"lj": "lj", "poly12": "poly12", "poly10": "poly10"}
RTOL, ATOL = 1e-10, 1e-12


@pytest.fixture(scope="module")
def vectors():
    return np.load(os.path.join(GOLDEN, "dim3_vectors.npz"))


def close(got, want, scale=None):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    err = np.abs(got - want)
    # components that vanish by symmetry are compared on the scale of the tensor they belong to
    floor = ATOL if scale is None else ATOL * max(1.0, scale)
    assert np.all(err <= floor + RTOL * np.abs(want)), "max abs err %.3e (max |want| %.3e)" % (err.max(), np.abs(want).max())


def load(vectors, case):
    z = {k: vectors["%s_%s" % (case, k)] for k in ("box", "pos", "diam", "vel", "mass", "disp", "grid", "virial", "elastic", "th", "meta")}
    model, kid, R, order, max_rcut, n_live = z["meta"]
    z.update(pot=POT_NAME[int(model)], kernel=KERNEL[int(kid)], R=float(R), order=int(order), max_rcut=float(max_rcut), n_live=int(n_live))
    return z


@pytest.mark.parametrize("case", CASES)
def test_oracle_dim3_matches_reference_calculators(orc, vectors, case):
    z = load(vectors, case)
    P = orc.potential(z["pot"])
    S = orc.System(z["box"], z["pos"], z["diam"], vel=z["vel"], mass=z["mass"], dim=3)
    assert S.n == z["n_live"] == len(z["diam"])              # periodic on all three axes: nobody dies
    assert S.max_rcut(P) == z["max_rcut"]
    cg = orc.cgfunc(z["kernel"], z["R"], z["order"])
    got = S.ik(P, cg, z["grid"], obs_mask=1)
    close(got["virial"], z["virial"], np.abs(z["virial"]).max())
    S.set_displacement(z["disp"])
    close(S.ik_elastic(P, cg, z["grid"]), z["elastic"], np.abs(z["elastic"]).max())
    th = S.thermo(P)
    want = z["th"]
    close(th["virial"], want[0:9], np.abs(want[0:9]).max())
    close(th["kinetic"], want[9:18], np.abs(want[9:18]).max())
    close(th["energy"], want[18])
    close(th["born"], want[19:100], np.abs(want[19:100]).max())
    close(S.thermo_elastic(P), want[100:109], np.abs(want[100:109]).max())
    # the brute-force search agrees with the bucketed one (27 images)
    b = S.ik(P, cg, z["grid"][:9], obs_mask=1, brute=True)
    close(b["virial"], got["virial"][:9], np.abs(z["virial"]).max())


def _gpu_setup(ctx, orc, z, f32=False):
    from pyglasstools_b200 import _capi  # noqa: F401
    P = orc.potential(z["pot"])
    ctx.set_potential(GPU_POT[z["pot"]], P.scaled_rcut, orc.pot_params(P))
    ctx.set_max_rcut(z["max_rcut"])
    dt = np.float32 if f32 else np.float64
    ctx.upload(z["pos"].astype(dt), z["diam"].astype(dt), z["box"], mass=z["mass"].astype(dt), vel=z["vel"].astype(dt), dim=3)
    ctx.build_cells()
    ctx.build_bonds()
    return P


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_dim3_fields_and_thermo_against_reference_vectors(orc, vectors, case):
    from pyglasstools_b200 import _capi
    z = load(vectors, case)
    ctx = _capi.Context(0)
    try:
        P = _gpu_setup(ctx, orc, z)
        assert ctx.counts()["n_live"] == z["n_live"]
        ctx.set_cg(z["kernel"], z["R"], z["order"])
        ctx.set_grid_points(z["grid"])
        ctx.set_displacement(z["disp"])
        ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_KINETIC | _capi.OBS_DENSITY | _capi.OBS_ELASTIC)
        vir = ctx.ik_download(_capi.FIELD_VIRIAL)
        assert vir.shape == (len(z["grid"]), 9)
        close(vir, z["virial"], np.abs(z["virial"]).max())
        close(ctx.ik_download(_capi.FIELD_ELASTIC), z["elastic"], np.abs(z["elastic"]).max())
        # local fields: the oracle (the reference's LocalField never sizes its tensors, tests/test_oracle_refloops.py)
        S = orc.System(z["box"], z["pos"], z["diam"], vel=z["vel"], mass=z["mass"], dim=3)
        want = S.ik(P, orc.cgfunc(z["kernel"], z["R"], z["order"]), z["grid"], obs_mask=2 | 4, kinetic_mode=0)
        close(ctx.ik_download(_capi.FIELD_KINETIC), want["kinetic"], np.abs(want["kinetic"]).max())
        close(ctx.ik_download(_capi.FIELD_DENSITY), want["density"])
        th = ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_KINETIC | _capi.OBS_ENERGY | _capi.OBS_BORN)
        w = z["th"]
        close(th["virial"], w[0:9], np.abs(w[0:9]).max())
        close(th["kinetic"], w[9:18], np.abs(w[9:18]).max())
        close(th["energy"], w[18])
        assert th["born"].shape == (81,)
        close(th["born"], w[19:100], np.abs(w[19:100]).max())
        close(ctx.thermo_elastic(), w[100:109], np.abs(w[100:109]).max())
        # observables not asked for come back as zeros
        only = ctx.thermo(_capi.OBS_ENERGY)
        assert not only["virial"].any() and not only["born"].any() and only["energy"] == th["energy"]
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_dim3_larger_frame_f32_cartesian_grid_shards_and_adaptive(orc):
    """A 4 096-particle float32 frame in a non-cubic box: Cartesian lattice (the ikcalculator's z = 0 plane) computed whole and
    in two grid shards, fixed-point and adaptive bond quadrature, against the oracle."""
    from pyglasstools_b200 import _capi
    rng = np.random.default_rng(77)
    n = 16
    L = np.array([16.0, 17.6, 14.4])
    ii, jj, kk = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    pos = ((np.stack([ii.ravel(), jj.ravel(), kk.ravel()], 1) + 0.5) / n - 0.5) * L + rng.uniform(-0.35, 0.35, (n**3, 3))
    pos = pos.astype(np.float32)
    pos[5] += np.float32(L[0])
    diam = rng.uniform(0.8, 1.25, n**3).astype(np.float32)
    vel = rng.standard_normal((n**3, 3)).astype(np.float32)
    P = orc.potential("poly12")
    S = orc.System(L, pos.astype(np.float64), diam.astype(np.float64), vel=vel.astype(np.float64), dim=3)
    ax = np.linspace(-L[0] / 2, L[0] / 2, 20)
    grid = np.zeros((400, 3)); grid[:, 0] = np.repeat(ax, 20); grid[:, 1] = np.tile(ax, 20)
    ctx = _capi.Context(0)
    try:
        ctx.set_potential("poly12", P.scaled_rcut, orc.pot_params(P))
        ctx.set_max_rcut(S.max_rcut(P))
        ctx.upload(pos, diam, L, vel=vel, dim=3)
        ctx.build_cells()
        ctx.build_bonds()
        assert ctx.counts()["n_live"] == n**3
        ctx.set_cg("octic", 2.0, 5)
        ctx.set_grid_cartesian(ax, ax)
        ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_KINETIC)
        want = S.ik(P, orc.cgfunc("octic", 2.0, 5), grid, obs_mask=1 | 2, kinetic_mode=0)
        full = ctx.ik_download(_capi.FIELD_VIRIAL)
        close(full, want["virial"], np.abs(want["virial"]).max())
        close(ctx.ik_download(_capi.FIELD_KINETIC), want["kinetic"], np.abs(want["kinetic"]).max())
        # grid shards (np.array_split of the i-major lattice): same bits as the whole grid
        for g0, g1 in ((0, 170), (170, 400)):
            ctx.set_shard(g0, g1)
            ctx.ik_compute(_capi.OBS_VIRIAL)
            assert np.array_equal(ctx.ik_download(_capi.FIELD_VIRIAL), full[g0:g1])
        ctx.set_shard(0, 400)
        th = ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_KINETIC)
        w = S.thermo(P, born=False)
        close(th["virial"], w["virial"], np.abs(w["virial"]).max())
        close(th["kinetic"], w["kinetic"], np.abs(w["kinetic"]).max())
        close(th["energy"], w["energy"])
        # adaptive (GSL QAG) bond quadrature on a few points
        ctx.set_cg_adaptive("octic", 2.0, 1e-9, 1e-9, 100)
        ctx.set_grid_points(grid[:40])
        ctx.ik_compute(_capi.OBS_VIRIAL)
        wa = S.ik(P, orc.cgfunc_adaptive("octic", 2.0, relerr=1e-9, abserr=1e-9, maxiter=100), grid[:40], obs_mask=1)
        close(ctx.ik_download(_capi.FIELD_VIRIAL), wa["virial"], np.abs(wa["virial"]).max())
        # 2-D only entry points say so
        with pytest.raises(_capi.PgtError):
            ctx.export_bonds()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_gpu_dim3_through_the_python_surface(orc, tmp_path, monkeypatch):
    """The example scripts' statements on a 3-D trajectory (utils.read_gsd(..., ndim=3)): ikcalculator + fieldlogger with
    3 x 3 component names and the x y z coordinate columns (io.py:112-128), thermo.calculator + logfile with a Born entry."""
    monkeypatch.chdir(tmp_path)
    import pyglasstools as pglass
    pglass.reset()
    try:
        import pyglasstools.io as io
        import pyglasstools.utils as utils
        import pyglasstools.potential as pot
        import pyglasstools.thermo as thermo
        import pyglasstools.irvingkirkwood as ik
        import pyglasstools.irvingkirkwood.cgfunc as cgfunc
        from pyglasstools import gsdio

        rng = np.random.default_rng(9)
        n, L = 8, 8.0
        frames = []
        for f in range(2):
            ii, jj, kk = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
            pos = (np.stack([ii.ravel(), jj.ravel(), kk.ravel()], 1) + 0.5 - n / 2 + rng.uniform(-0.3, 0.3, (n**3, 3))).astype(np.float32)
            diam = rng.uniform(0.8, 1.25, n**3).astype(np.float32)
            frames.append(dict(box=np.array([L, L, L, 0, 0, 0], np.float32), position=pos, diameter=diam, dimensions=3))
        gsdio.write_trajectory("traj3.gsd", frames)
        poly12 = pot.polydisperse12(v0=1.0, eps=0.2, rcut=1.25)                                # noqa: F841
        mycgfunc = cgfunc.octic(rcut=1.5, mode="fixedpoint", order=5)
        mysystem = utils.read_gsd(filename="traj3.gsd", checkpointfile="restart.log", ndim=3)  # noqa: F841
        assert pglass.get_sysdata().pysimbox.dim == 3
        calc = ik.ikcalculator(cgfunc=mycgfunc, dx=1.0)
        names = ['virialstress_00', 'virialstress_22', 'virialstress_12']
        flog = io.fieldlogger(keyword="dump3", names=names, solver=calc)
        tcalc = thermo.calculator()
        tlog = io.logfile(filename="thermo3.log", names=['virialstress_22', 'virialstress_02', 'potentialenergy', 'borntensor_0122'], solver=tcalc)
        pglass.analyze(frame_list=range(2))
        flog.file.flush()
        del tlog

        P, cg = orc.potential("poly12"), orc.cgfunc("octic", 1.5, 5)
        grid = np.asarray(calc.gridpoints)
        assert grid.shape == (64, 3) and not grid[:, 2].any()
        S0 = orc.System(frames[0]["box"][:3], frames[0]["position"], frames[0]["diameter"], dim=3)
        mr = S0.max_rcut(P)                                      # frozen at the calculator's construction (Calculator.h:37-46)
        S = orc.System(frames[1]["box"][:3], frames[1]["position"], frames[1]["diameter"], dim=3)
        want = S.ik(P, cg, grid, max_rcut=mr)["virial"]
        got = flog.field_obs['virialstress'].values()
        assert got.shape == (64, 9)
        close(got, want, np.abs(want).max())
        lines = open("dump3.xyz").read().split("\n")
        assert lines[0] == "64 " and lines[1].startswith("#Frame 0")
        last = lines[66 + 2:66 + 2 + 64]                          # second frame's rows
        for g in (0, 17, 63):
            cols = last[g].split()
            assert len(cols) == 1 + 3 + 3                         # "1 x y z" + three components
            np.testing.assert_allclose([float(c) for c in cols[1:4]], grid[g], rtol=1e-5, atol=1e-12)
            np.testing.assert_allclose([float(c) for c in cols[4:]], [want[g][0], want[g][8], want[g][1 + 3 * 2]], rtol=2e-5, atol=1e-9)
        th = np.loadtxt("thermo3.log", skiprows=1)
        assert th.shape == (2, 5)
        w = S.thermo(P, max_rcut=mr)
        born_0122 = w["born"][2 + 3 * (2 + 3 * (1 + 3 * 0))]      # io.py: 'abcd' -> d + dim (c + dim (b + dim a))
        np.testing.assert_allclose(th[1, 1:], [w["virial"][8], w["virial"][0 + 3 * 2], w["energy"], born_0122], rtol=0, atol=1e-6)
    finally:
        pglass.reset()
