"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on the same inputs.

Tolerances are BASELINE.json's: bit-exact neighbour / bond sets; stress and energy within 1e-10
relative (1e-12 absolute for near-zero components).
"""
import numpy as np
import pytest

from pyglasstools_b200 import _capi, synth

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-10, 1e-12


WORST = {"rel": 0.0}       # largest per-component relative error seen among components above the absolute floor


def assert_close(a, b, rtol=RTOL, atol=ATOL):
    """BASELINE.json's tolerance, applied PER COMPONENT: |got - want| <= 1e-12 + 1e-10 |want| element-wise."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    err = np.abs(a - b)
    bad = err > (atol + rtol * np.abs(b))
    big = np.abs(b) > 1e3 * atol
    rel = float((err[big] / np.abs(b[big])).max()) if big.any() else 0.0
    WORST["rel"] = max(WORST["rel"], rel)
    print("max per-component relative error %.3e (max abs %.3e)" % (rel, float(err.max()) if err.size else 0.0))
    assert not bad.any(), "max abs err %.3e, max per-component rel err %.3e, %d entries out of tolerance" % (
        err.max(), rel, int(bad.sum()))


def setup_ctx(ctx, orc, P, name, pos, diam, box, vel=None, mass=None):
    S = orc.System(box, pos, diam, vel=vel, mass=mass)
    ctx.set_potential(name, P.scaled_rcut, orc.pot_params(P))
    ctx.set_max_rcut(S.max_rcut(P))
    ctx.upload(pos, diam, box, mass=mass, vel=vel)
    ctx.build_cells()
    return S


GPU_NAME = {"poly12": "poly12", "poly18": "poly18", "poly10": "poly10", "poly106": "poly106",
            "polyyukawa": "polyyukawa", "polylj": "polylj", "lj": "lj", "fslj": "forceshifted-lj"}


def pairs_key(i, j):
    return set(zip(i.tolist(), j.tolist()))


@pytest.mark.parametrize("frame", [0, 180, 199])
def test_fixture_neighbor_and_bond_sets_bit_exact(gpu_ctx, orc, fixture_frames, frame):
    P = orc.potential("poly12")
    pos, diam, box = fixture_frames.position(frame), fixture_frames.diameter, fixture_frames.box
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    oi, oj, orij, ofl = S.neighbor_pairs(P)
    gi, gj = gpu_ctx.export_neighbors()
    assert len(gi) == len(oi)
    assert pairs_key(gi, gj) == pairs_key(oi, oj)
    bi, bj, brij, bf = gpu_ctx.export_bonds()
    m = ofl.astype(bool)
    assert len(bi) == int(m.sum())
    assert pairs_key(bi, bj) == pairs_key(oi[m], oj[m])
    # r_ij and f of every bond, matched by (i,j)
    ref = {(a, b): (r, k) for a, b, r, k in zip(oi[m].tolist(), oj[m].tolist(), orij[m], range(int(m.sum())))}
    fd = orc.pair_table(P, orij[m], S_diam(S, oi[m], diam), S_diam(S, oj[m], diam))["fdivr"]
    for a, b, r, f in zip(bi.tolist(), bj.tolist(), brij, bf):
        rr, k = ref[(a, b)]
        assert r[0] == rr[0] and r[1] == rr[1]          # bit-exact image-resolved separation
        assert f == fd[k]                               # bit-exact force (same operation order)


def S_diam(S, idx, diam):
    return np.asarray(diam, np.float64)[idx]


@pytest.mark.parametrize("frame", [180, 190, 199])
def test_fixture_ik_virial_field(gpu_ctx, orc, fixture_frames, frame):
    """cfg1: example_scripts/irvingkirkwood on the shipped frames (octic R=2, GL-5, dx=0.25)."""
    P = orc.potential("poly12")
    cg = orc.cgfunc("octic", 2.0, 5)
    pos, diam, box = fixture_frames.position(frame), fixture_frames.diameter, fixture_frames.box
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    nmax = int(L / 0.25)
    axis = synth.linspace_axis(L, nmax)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_DENSITY)
    got = gpu_ctx.ik_download(_capi.FIELD_VIRIAL)
    grid, n2 = orc.cartesian_grid(L, 0.25)
    assert n2 == nmax == 127
    want = S.ik(P, cg, grid, obs_mask=1 | 4)
    assert_close(got, want["virial"])
    assert_close(gpu_ctx.ik_download(_capi.FIELD_DENSITY), want["density"])


@pytest.mark.parametrize("frame", [0, 57, 199])
def test_fixture_thermo(gpu_ctx, orc, fixture_frames, frame):
    P = orc.potential("poly12")
    pos, diam, box = fixture_frames.position(frame), fixture_frames.diameter, fixture_frames.box
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    got = gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_BORN | _capi.OBS_KINETIC)
    want = S.thermo(P)
    assert_close(got["virial"], want["virial"])
    assert_close(got["energy"], want["energy"])
    assert_close(got["born"], want["born"])
    assert np.all(got["kinetic"] == 0)


def test_born_known_answer_on_gpu(gpu_ctx, orc, fixture_frames):
    """The reference README's only known-answer value (README.md:149-172), computed on the GPU."""
    P = orc.potential("poly12")
    b0101, bdiff = [], []
    D = 2
    idx = lambda i, j, k, l: l + D * (k + D * (j + D * i))      # io.py:78
    rd = lambda v: float("%f" % v)                              # std::to_string, ThermoProperty.h:196
    for f in range(200):
        setup_ctx(gpu_ctx, orc, P, "poly12", fixture_frames.position(f), fixture_frames.diameter, fixture_frames.box)
        r = gpu_ctx.thermo(_capi.OBS_BORN)
        b0101.append(rd(r["born"][idx(0, 1, 0, 1)]))
        bdiff.append(rd(r["born"][idx(0, 0, 0, 0)]) - rd(r["born"][idx(0, 0, 1, 1)]))
    assert abs(np.mean(b0101) - 14.89655895) < 5e-9
    assert abs(0.5 * np.mean(bdiff) - 14.8769594875) < 5e-9


@pytest.mark.parametrize("model", ["poly18", "poly10", "poly106", "polylj", "polyyukawa", "lj", "fslj"])
def test_other_potentials_ik_and_thermo(gpu_ctx, orc, model):
    fr = synth.synth_frame(32, 77)
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential(model)
    S = setup_ctx(gpu_ctx, orc, P, GPU_NAME[model], pos, diam, box)
    gpu_ctx.build_bonds()
    oi, oj, orij, ofl = S.neighbor_pairs(P)
    bi, bj, _, _ = gpu_ctx.export_bonds()
    m = ofl.astype(bool)
    assert pairs_key(bi, bj) == pairs_key(oi[m], oj[m])
    got = gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_BORN)
    want = S.thermo(P)
    assert_close(got["virial"], want["virial"])
    assert_close(got["energy"], want["energy"])
    assert_close(got["born"], want["born"])
    cg = orc.cgfunc("octic", 2.0, 5)
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, 40)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    grid = np.zeros((1600, 3)); grid[:, 0] = np.repeat(axis, 40); grid[:, 1] = np.tile(axis, 40)
    wantf = S.ik(P, cg, grid)["virial"]
    assert_close(gpu_ctx.ik_download(), wantf)


@pytest.mark.parametrize("kernel,order,R", [("octic", 1, 2.0), ("octic", 2, 2.5), ("octic", 8, 2.0), ("octic", 20, 3.0),
                                             ("octic", 7, 4.0), ("mollifier", 5, 2.0), ("mollifier", 12, 2.5),
                                             ("rect", 5, 2.0)])
def test_cg_kernels_and_orders(gpu_ctx, orc, kernel, order, R):
    fr = synth.synth_frame(24, 5)
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    cg = orc.cgfunc(kernel, R, order)
    gpu_ctx.set_cg(kernel, R, order)
    rng = np.random.default_rng(3)
    L = float(box[0])
    pts = np.zeros((500, 3)); pts[:, :2] = rng.uniform(-L / 2, L / 2, (500, 2))
    gpu_ctx.set_grid_points(pts)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_DENSITY)
    want = S.ik(P, cg, pts, obs_mask=1 | 4)
    got = gpu_ctx.ik_download()
    # rect is discontinuous at the disc edge: the device tests its nodes in the reference's operation order
    # (CoarseGrainFunction.h:77-88), so no grid point may differ
    assert_close(got, want["virial"])
    assert_close(gpu_ctx.ik_download(_capi.FIELD_DENSITY), want["density"])


def test_kinetic_field_and_thermo_kinetic(gpu_ctx, orc):
    fr = synth.synth_frame(24, 11, velocities=True)
    pos, diam, box, vel = fr["position"], fr["diameter"], fr["box"], fr["velocity"]
    mass = (1.0 + np.arange(len(diam)) % 3).astype(np.float32)
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box, vel=vel, mass=mass)
    gpu_ctx.build_bonds()
    cg = orc.cgfunc("octic", 2.0, 5)
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, 30)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_KINETIC)
    grid = np.zeros((900, 3)); grid[:, 0] = np.repeat(axis, 30); grid[:, 1] = np.tile(axis, 30)
    want = S.ik(P, cg, grid, obs_mask=1 | 2, kinetic_mode=0)
    assert_close(gpu_ctx.ik_download(_capi.FIELD_KINETIC), want["kinetic"])
    assert_close(gpu_ctx.ik_download(_capi.FIELD_VIRIAL), want["virial"])
    t = gpu_ctx.thermo(_capi.OBS_KINETIC)
    assert_close(t["kinetic"], S.thermo(P)["kinetic"])


def test_edge_cases_seam_disc_edge_and_outside_points(gpu_ctx, orc):
    """Particles on the periodic seam, exactly on the disc edge (||r-x|| == R), exactly at the pair
    cut-off, a deleted particle (z outside the slab), unwrapped inputs, grid points at +-L/2."""
    L = 16.0
    box = np.array([L, L, 1.0])
    pos = []
    rng = np.random.default_rng(9)
    for i in range(8):
        for j in range(8):
            pos.append([-8 + 2 * i + 0.25 + 0.5 * rng.random(), -8 + 2 * j + 0.25 + 0.5 * rng.random(), 0.0])
    pos += [[-8.0, 0.0, 0.0], [7.99999, 0.5, 0.0], [0.0, -8.0, 0.0], [0.3, 7.9999, 0.0]]      # seam
    pos += [[2.0, 0.0, 0.0], [0.0, -2.0, 0.0]]                  # exactly R=2 from the grid point (0,0)
    pos += [[24.5, 1.0, 0.0], [-17.0, -9.5, 0.0]]               # need wrapping by one / two box lengths
    pos += [[1.0, 1.0, 0.75]]                                   # outside the z slab: deleted by Aboria
    pos = np.array(pos, np.float64)
    diam = np.full(len(pos), 1.0); diam[::3] = 1.25; diam[1::5] = 0.8
    # a pair exactly at the cut-off: sigma = 1 (d=1,1), rcut = 1.25 -> r = 1.25 gives rsq == rcut^2 sigma^2 (strict <: no bond)
    pos = np.vstack([pos, [[5.0, 5.0, 0.0], [6.25, 5.0, 0.0]]])
    diam = np.concatenate([diam, [1.0, 1.0]])
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    assert gpu_ctx.counts()["n_live"] == S.n == len(pos) - 1
    wp, ids = S.wrapped()
    gxy, alive = gpu_ctx.export_wrapped(len(pos))
    assert alive.sum() == S.n and not alive[len(pos) - 3]
    assert np.array_equal(gxy[ids], wp[:, :2])
    gpu_ctx.build_bonds()
    oi, oj, orij, ofl = S.neighbor_pairs(P, brute=True)
    gi, gj = gpu_ctx.export_neighbors()
    assert pairs_key(gi, gj) == pairs_key(oi, oj)
    bi, bj, _, _ = gpu_ctx.export_bonds()
    m = ofl.astype(bool)
    assert pairs_key(bi, bj) == pairs_key(oi[m], oj[m])
    n = len(pos)
    assert (n - 2, n - 1) in pairs_key(gi, gj) and (n - 2, n - 1) not in pairs_key(bi, bj)
    cg = orc.cgfunc("octic", 2.0, 5)
    gpu_ctx.set_cg("octic", 2.0, 5)
    pts = np.array([[0, 0, 0], [8.0, 8.0, 0], [-8.0, -8.0, 0], [8.0, 0.1, 0], [-8.0, 3.0, 0], [7.5, -7.5, 0],
                    [0.0, 8.0, 0], [3.0, 3.0, 0], [12.0, 0.0, 0], [-30.0, 2.0, 0]], np.float64)
    gpu_ctx.set_grid_points(pts)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_DENSITY)
    want = S.ik(P, cg, pts, obs_mask=1 | 4, brute=True, counts=True)
    assert want["counts"][0, 0] >= 2       # the two edge particles are inside (inclusive <=)
    assert_close(gpu_ctx.ik_download(), want["virial"])
    assert_close(gpu_ctx.ik_download(_capi.FIELD_DENSITY), want["density"])
    # the periodic duplicates: the field at +L/2 equals the field at -L/2
    f = gpu_ctx.ik_download()
    assert_close(f[1], f[2])


def test_empty_and_tiny_inputs(gpu_ctx, orc):
    P = orc.potential("poly12")
    box = np.array([12.0, 12.0, 1.0])
    pos = np.array([[0.0, 0.0, 0.0], [0.9, 0.1, 0.0]])
    diam = np.array([1.0, 1.1])
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    assert gpu_ctx.counts()["n_halfbonds"] == 2
    gpu_ctx.set_cg("octic", 2.0, 5)
    gpu_ctx.set_grid_points(np.zeros((0, 3)))
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    assert gpu_ctx.ik_download().shape == (0, 4)
    pts = np.array([[0.3, 0.0, 0.0], [5.0, 5.0, 0.0]])
    gpu_ctx.set_grid_points(pts)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    cg = orc.cgfunc("octic", 2.0, 5)
    want = S.ik(P, cg, pts)["virial"]
    got = gpu_ctx.ik_download()
    assert_close(got, want)
    assert np.all(got[1] == 0)
    t = gpu_ctx.thermo()
    assert_close(t["energy"], S.thermo(P)["energy"])
    # a frame with no force-carrying pair at all
    pos2 = np.array([[0.0, 0.0, 0.0], [4.0, 4.0, 0.0]])
    setup_ctx(gpu_ctx, orc, P, "poly12", pos2, diam, box)
    gpu_ctx.build_bonds()
    assert gpu_ctx.counts()["n_halfbonds"] == 0
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    assert np.all(gpu_ctx.ik_download() == 0)


@pytest.mark.parametrize("variant", [2, 7])
def test_sharding_is_bitwise_independent(gpu_ctx, orc, fixture_frames, variant):
    """np.array_split shards (irvingkirkwood/__init__.py:53-60) reproduce the unsharded field bit for bit,
    with either gather kernel."""
    gpu_ctx.set_option("ik_variant", variant)
    P = orc.potential("poly12")
    pos, diam, box = fixture_frames.position(185), fixture_frames.diameter, fixture_frames.box
    setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, 127)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    full = gpu_ctx.ik_download()
    G = 127 * 127
    for nshard in (2, 5, 8):
        sizes = [len(c) for c in np.array_split(np.arange(G), nshard)]
        parts, g0 = [], 0
        for s in sizes:
            gpu_ctx.set_shard(g0, g0 + s)
            gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
            parts.append(gpu_ctx.ik_download())
            g0 += s
        assert np.array_equal(np.vstack(parts), full)
    gpu_ctx.set_shard(0, G)
    gpu_ctx.set_option("ik_variant", 0)


def test_bond_list_single_search_equals_two_search(gpu_ctx, orc, fixture_frames):
    """The one-search bond builder (neighbours kept, dx re-formed from the same operands) writes the same records,
    bit for bit and in the same order, as the count + fill pair; so does the path taken when a particle has more
    bonds than the kept list holds (poly106, rc = 3: ~28 bonds per particle)."""
    for model in ("poly12", "poly106"):
        P = orc.potential(model)
        pos, diam, box = fixture_frames.position(33), fixture_frames.diameter, fixture_frames.box
        got = []
        for single, flat in ((1, 1), (1, 0), (0, 1)):
            gpu_ctx.set_option("bond_single_search", single)
            gpu_ctx.set_option("bond_fill_flat", flat)       # warp-flattened writer vs thread per particle
            setup_ctx(gpu_ctx, orc, P, GPU_NAME[model], pos, diam, box)
            gpu_ctx.build_bonds()
            got.append(gpu_ctx.export_bonds())
        gpu_ctx.set_option("bond_single_search", 1)
        gpu_ctx.set_option("bond_fill_flat", 1)
        for other in got[1:]:
            for a, b in zip(got[0], other):
                assert np.array_equal(a, b)
        assert len(got[0][0]) > (20000 if model == "poly106" else 5000)


@pytest.mark.parametrize("variant", [0, 2, 7])
def test_gather_kernel_variants(orc, fixture_frames, variant, monkeypatch):
    """Every gather formulation (two-phase point lists, quad gather over bulk-staged tiles) against the oracle."""
    monkeypatch.setenv("PGT_IK_VARIANT", str(variant))
    ctx = _capi.Context(0)
    try:
        P = orc.potential("poly12")
        cg = orc.cgfunc("octic", 2.0, 5)
        pos, diam, box = fixture_frames.position(181), fixture_frames.diameter, fixture_frames.box
        S = setup_ctx(ctx, orc, P, "poly12", pos, diam, box)
        ctx.build_bonds()
        ctx.set_cg("octic", 2.0, 5)
        L = float(box[0])
        axis = synth.linspace_axis(L, 60)
        ctx.set_grid_cartesian(axis, axis)
        ctx.ik_compute(_capi.OBS_VIRIAL)
        grid = np.zeros((3600, 3)); grid[:, 0] = np.repeat(axis, 60); grid[:, 1] = np.tile(axis, 60)
        want = S.ik(P, cg, grid)["virial"]
        assert_close(ctx.ik_download(), want)
    finally:
        ctx.close()


def test_synthetic_large_frame_properties(gpu_ctx, orc):
    """N = 65536 (cfg2 size): full thermo parity, IK parity on a strided subsample of a 256^2 grid, and
    the size-independent checks: bond symmetry (every (i,j) has its (j,i)) and the IK field summed
    against the oracle's own sum."""
    fr = synth.synth_frame(256, synth.SEEDS["cfg2"])
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    bi, bj, brij, bf = gpu_ctx.export_bonds()
    fwd = dict(zip(zip(bi.tolist(), bj.tolist()), range(len(bi))))
    for (a, b), k in list(fwd.items())[:: max(1, len(fwd) // 5000)]:
        k2 = fwd[(b, a)]
        assert bf[k] == bf[k2] and np.array_equal(brij[k], -brij[k2])
    got = gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY)
    want = S.thermo(P, born=False)
    assert_close(got["virial"], want["virial"])
    assert_close(got["energy"], want["energy"])
    cg = orc.cgfunc("octic", 2.0, 5)
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, 256)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    full = gpu_ctx.ik_download().reshape(256, 256, 4)
    ii = np.arange(0, 256, 5)
    grid = np.zeros((len(ii) ** 2, 3)); grid[:, 0] = np.repeat(axis[ii], len(ii)); grid[:, 1] = np.tile(axis[ii], len(ii))
    wantf = S.ik(P, cg, grid)["virial"]
    gotf = full[np.ix_(ii, ii)].reshape(-1, 4)
    assert_close(gotf, wantf)
    assert np.array_equal(full[0], full[-1]) or np.allclose(full[0], full[-1], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("kernel,order,R,nmax,model", [("octic", 5, 2.0, 97, "poly12"), ("octic", 5, 4.0, 64, "poly10"),
                                                       ("octic", 3, 1.0, 200, "poly12"), ("mollifier", 5, 2.5, 50, "poly12"),
                                                       ("octic", 8, 3.0, 31, "poly106")])
def test_quad_gather_on_cartesian_grids(gpu_ctx, orc, kernel, order, R, nmax, model):
    """The quad gather (variant 7: unique bonds, weights 1 and 1/2) on Cartesian grids whose tiles are interior, on the
    periodic seam and partial (nmax not a multiple of the tile), for several radii / quadratures / potentials:
    against the oracle on a subsample and against the two-phase kernel everywhere."""
    fr = synth.synth_frame(48, 17)
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential(model)
    S = setup_ctx(gpu_ctx, orc, P, GPU_NAME[model], pos, diam, box)
    gpu_ctx.build_bonds()
    cg = orc.cgfunc(kernel, R, order)
    gpu_ctx.set_cg(kernel, R, order)
    L = float(box[0])
    axis = synth.linspace_axis(L, nmax)
    gpu_ctx.set_grid_cartesian(axis, axis)
    fields = {}
    for variant in (2, 7):
        gpu_ctx.set_option("ik_variant", variant)
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        fields[variant] = gpu_ctx.ik_download()
    gpu_ctx.set_option("ik_variant", 0)
    assert_close(fields[7], fields[2])
    ii = np.unique(np.concatenate([np.arange(0, nmax, 7), [nmax - 1]]))
    grid = np.zeros((len(ii) ** 2, 3)); grid[:, 0] = np.repeat(axis[ii], len(ii)); grid[:, 1] = np.tile(axis[ii], len(ii))
    want = S.ik(P, cg, grid)["virial"]
    got = fields[7].reshape(nmax, nmax, 4)[np.ix_(ii, ii)].reshape(-1, 4)
    assert_close(got, want)


@pytest.mark.parametrize("kernel", ["octic", "mollifier"])
def test_adaptive_bond_quadrature_field(gpu_ctx, orc, kernel):
    """cgfunc.<kernel>(mode="adaptive"): the device QAG / GK15 field against the oracle's restatement of
    gsl_integration_qag (same bisection rule, so far inside the tolerance the user asked for), against the fixed
    1024-point rule within that tolerance, and the failure exit the reference turns into an abort."""
    fr = synth.synth_frame(24, 9)
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    rng = np.random.default_rng(5)
    L = float(box[0])
    pts = np.zeros((300, 3)); pts[:, :2] = rng.uniform(-L / 2, L / 2, (300, 2))
    gpu_ctx.set_grid_points(pts)
    gpu_ctx.set_cg_adaptive(kernel, 2.0, relerr=1e-9, abserr=1e-11, maxiter=100)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_DENSITY)
    got = gpu_ctx.ik_download()
    rho = gpu_ctx.ik_download(_capi.FIELD_DENSITY)
    want = S.ik(P, orc.cgfunc_adaptive(kernel, 2.0, 1e-9, 1e-11, 100), pts, obs_mask=1 | 4)
    scale = np.abs(want["virial"]).max(axis=1, keepdims=True)
    assert_close(got, want["virial"])
    assert_close(rho, want["density"])
    fixed = S.ik(P, orc.cgfunc(kernel, 2.0, 1024), pts)["virial"]
    assert np.all(np.abs(got - fixed) <= 1e-6 * scale + 1e-8)
    # Cartesian grids take the same kernel
    axis = synth.linspace_axis(L, 20)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    grid = np.zeros((400, 3)); grid[:, 0] = np.repeat(axis, 20); grid[:, 1] = np.tile(axis, 20)
    wantg = S.ik(P, orc.cgfunc_adaptive(kernel, 2.0, 1e-9, 1e-11, 100), grid)["virial"]
    assert_close(gpu_ctx.ik_download(), wantg)
    # two intervals cannot reach 1e-13: GSL_EMAXITER -> the reference aborts, we return an error
    gpu_ctx.set_cg_adaptive(kernel, 2.0, relerr=1e-13, abserr=1e-300, maxiter=2)
    with pytest.raises(_capi.PgtError):
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    gpu_ctx.set_cg("octic", 2.0, 5)


@pytest.mark.parametrize("model", ["poly12", "polylj"])
def test_elastic_virial_stress_global_and_field(gpu_ctx, orc, model):
    """ElasticVirialStress (ListofCommonObservables.h:86-123): F1 (x) (u+r) - F (x) r with per-particle displacements u,
    as the global sum (ThermoCalculator, r = +dx) and as the IK field (IrvingKirkwood hands the functor r = -dx: the
    quirk is reproduced).  Zero displacement gives exactly zero; pairs beyond their cut-off at r but inside it at
    r + u are counted (the kernels search max_rcut, they do not read the bond list)."""
    fr = synth.synth_frame(24, 21)
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential(model)
    S = setup_ctx(gpu_ctx, orc, P, GPU_NAME[model], pos, diam, box)
    gpu_ctx.build_bonds()
    assert np.all(gpu_ctx.thermo_elastic() == 0.0)                      # ParticleSystem.cc:28: displacement = 0
    rng = np.random.default_rng(8)
    disp = np.zeros((len(diam), 3)); disp[:, :2] = rng.normal(0.0, 0.08, (len(diam), 2))
    gpu_ctx.set_displacement(disp)
    S.set_displacement(disp)
    want = S.thermo_elastic(P)
    assert_close(gpu_ctx.thermo_elastic(), want)
    L = float(box[0])
    pts = np.zeros((200, 3)); pts[:, :2] = rng.uniform(-L / 2, L / 2, (200, 2))
    gpu_ctx.set_grid_points(pts)
    for cgname, args in (("fixed", ("octic", 2.0, 5)), ("adaptive", ("octic", 2.0, 1e-9, 1e-11, 100))):
        if cgname == "fixed":
            gpu_ctx.set_cg(*args); cg = orc.cgfunc(*args)
        else:
            gpu_ctx.set_cg_adaptive(*args); cg = orc.cgfunc_adaptive(*args)
        gpu_ctx.ik_compute(_capi.OBS_ELASTIC | _capi.OBS_VIRIAL)
        got = gpu_ctx.ik_download(_capi.FIELD_ELASTIC)
        wantf = S.ik_elastic(P, cg, pts)
        assert_close(got, wantf)
        wantv = S.ik(P, cg, pts)["virial"]
        assert_close(gpu_ctx.ik_download(_capi.FIELD_VIRIAL), wantv)
    gpu_ctx.set_cg("octic", 2.0, 5)
    # a new frame forgets the displacements
    setup_ctx(gpu_ctx, orc, P, GPU_NAME[model], pos, diam, box)
    assert np.all(gpu_ctx.thermo_elastic() == 0.0)


@pytest.mark.parametrize("n,nsample", [(1024, 300), (2048, 120)])
def test_full_size_frames_properties_and_sampled_oracle(gpu_ctx, orc, n, nsample):
    """BASELINE configs[2] / configs[4] sizes (N = 1 048 576 on 1024^2, N = 4 194 304 on 2048^2): the oracle on a random
    sample of grid points, and the size-independent checks over the whole field: the two gather formulations agree
    everywhere, the duplicated end rows / columns of the np.linspace grid agree, a 3-way np.array_split sharding
    reproduces the field bit for bit, half-bonds come in pairs, and the thermo virial equals the sum over the exported
    bond list."""
    fr = synth.synth_frame(n, synth.SEEDS["cfg3" if n == 1024 else "cfg5"])
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    nhb = gpu_ctx.counts()["n_halfbonds"]
    assert nhb % 2 == 0 and nhb > 3 * n * n
    gpu_ctx.set_cg("octic", 2.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, n)
    gpu_ctx.set_grid_cartesian(axis, axis)
    fields = {}
    for variant in (2, 7):
        gpu_ctx.set_option("ik_variant", variant)
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        fields[variant] = gpu_ctx.ik_download()
    full = fields[7].reshape(n, n, 4)
    assert_close(fields[7], fields[2])
    # x = -L/2 and x = +L/2 are the same point of the periodic box (irvingkirkwood/__init__.py:45-53)
    assert_close(full[0], full[-1])
    assert_close(full[:, 0], full[:, -1])
    # sharding: bit for bit, with the automatic kernel choice
    gpu_ctx.set_option("ik_variant", 0)
    G = n * n
    parts, g0 = [], 0
    for s in [len(c) for c in np.array_split(np.arange(G), 3)]:
        gpu_ctx.set_shard(g0, g0 + s)
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        parts.append(gpu_ctx.ik_download())
        g0 += s
    gpu_ctx.set_shard(0, G)
    assert np.array_equal(np.vstack(parts), fields[7])      # automatic = the quad gather on this grid
    # the oracle on a random sample of grid points (incl. the seam rows)
    rng = np.random.default_rng(n)
    ii = np.concatenate([rng.integers(0, n, nsample), [0, n - 1, 0, n - 1]])
    jj = np.concatenate([rng.integers(0, n, nsample), [0, n - 1, n - 1, 0]])
    grid = np.zeros((len(ii), 3)); grid[:, 0] = axis[ii]; grid[:, 1] = axis[jj]
    want = S.ik(P, orc.cgfunc("octic", 2.0, 5), grid)["virial"]
    assert_close(full[ii, jj], want)
    # thermo: device pair sums vs the oracle's serial loop, and vs the bond list (each unordered pair appears twice)
    t = gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY)
    if n == 1024:
        w = S.thermo(P, born=False)
        assert_close(t["virial"], w["virial"])
        assert_close(t["energy"], w["energy"])
    bi, bj, brij, bf = gpu_ctx.export_bonds()
    vol = float(box[0]) * float(box[1])
    vxx = 0.5 * np.sum(bf * brij[:, 0] * brij[:, 0]) / vol
    vyy = 0.5 * np.sum(bf * brij[:, 1] * brij[:, 1]) / vol
    assert abs(vxx - t["virial"][0]) <= 1e-9 * abs(vxx) and abs(vyy - t["virial"][3]) <= 1e-9 * abs(vyy)
    assert np.array_equal(np.sort(bi.astype(np.int64) * (n * n) + bj), np.sort(bj.astype(np.int64) * (n * n) + bi))


def test_full_size_cfg4_other_potential_wide_kernel(gpu_ctx, orc):
    """BASELINE configs[3] at full size: polydisperse10 (q = 3, rcut 1.48), octic R = 4, GL-5, N = 1 048 576 on a 1024^2
    grid.  The oracle on a random sample of grid points (with the seam corners); over the whole field the two gather
    formulations agree, the duplicated end rows / columns agree and a 4-way np.array_split sharding (slab-local cell and
    bond lists) reproduces the field bit for bit."""
    n = 1024
    fr = synth.synth_frame(n, synth.SEEDS["cfg4"])
    pos, diam, box = fr["position"], fr["diameter"], fr["box"]
    P = orc.potential("poly10")
    S = setup_ctx(gpu_ctx, orc, P, "poly10", pos, diam, box)
    gpu_ctx.build_bonds()
    gpu_ctx.set_cg("octic", 4.0, 5)
    L = float(box[0])
    axis = synth.linspace_axis(L, n)
    gpu_ctx.set_grid_cartesian(axis, axis)
    fields = {}
    for variant in (2, 7):
        gpu_ctx.set_option("ik_variant", variant)
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        fields[variant] = gpu_ctx.ik_download()
    gpu_ctx.set_option("ik_variant", 0)
    full = fields[7].reshape(n, n, 4)
    assert_close(fields[7], fields[2])
    assert_close(full[0], full[-1])
    assert_close(full[:, 0], full[:, -1])
    rng = np.random.default_rng(44)
    ii = np.concatenate([rng.integers(0, n, 150), [0, n - 1, 0, n - 1]])
    jj = np.concatenate([rng.integers(0, n, 150), [0, n - 1, n - 1, 0]])
    grid = np.zeros((len(ii), 3)); grid[:, 0] = axis[ii]; grid[:, 1] = axis[jj]
    want = S.ik(P, orc.cgfunc("octic", 4.0, 5), grid)["virial"]
    assert_close(full[ii, jj], want)
    # sharded: every shard re-builds its own slab of the frame
    G = n * n
    parts, g0, kept = [], 0, []
    for s in [len(c) for c in np.array_split(np.arange(G), 4)]:
        gpu_ctx.set_shard(g0, g0 + s)
        gpu_ctx.upload(pos, diam, box)
        gpu_ctx.build_cells()
        gpu_ctx.build_bonds()
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        parts.append(gpu_ctx.ik_download())
        kept.append(gpu_ctx.counts()["n_live"])
        g0 += s
    gpu_ctx.set_shard(0, G)
    assert max(kept) < 0.4 * n * n
    assert np.array_equal(np.vstack(parts), fields[7])


def test_deferred_counts_same_bits_and_late_errors(orc, fixture_frames):
    """Frames of the same shape as one already counted do not wait for their own particle / half-bond counts (option
    "async_counts", default on): same fields bit for bit as with exact counting; a frame whose bond count outgrows the
    margin is reported at the next synchronisation point and the frame can be submitted again."""
    diam, box = fixture_frames.diameter, fixture_frames.box
    P = orc.potential("poly12")
    L = float(box[0])
    axis = synth.linspace_axis(L, 64)
    fields = {}
    for mode in (1, 0):
        ctx = _capi.Context(0)
        ctx.set_option("async_counts", mode)
        ctx.set_potential("poly12", P.scaled_rcut, orc.pot_params(P))
        ctx.set_cg("octic", 2.0, 5)
        ctx.set_max_rcut(orc.System(box, fixture_frames.position(0), diam).max_rcut(P))
        ctx.set_grid_cartesian(axis, axis)
        out = []
        pinned = [_capi.PinnedArray((64 * 64, 4), np.float64) for _ in range(6)]
        for k, f in enumerate((0, 40, 80, 120, 160, 199)):
            ctx.upload(fixture_frames.position(f), diam, box)
            ctx.build_cells()
            ctx.build_bonds()
            ctx.ik_compute(_capi.OBS_VIRIAL)
            ctx.ik_download_async(_capi.FIELD_VIRIAL, pinned[k].array)
        ctx.sync()
        fields[mode] = [p.array.copy() for p in pinned]
        counts = ctx.counts()
        assert counts["n_live"] == len(diam) and counts["n_halfbonds"] > 0
        if mode == 1:
            # a frame with many more bonds than the margin allows: the error arrives late, then the frame goes through
            big = (diam * 2.5).astype(np.float32)               # every neighbour within max_rcut bonds: beyond the margin (and 16 slots)
            ctx.upload(fixture_frames.position(199), big, box)
            ctx.build_cells()
            ctx.build_bonds()
            ctx.ik_compute(_capi.OBS_VIRIAL)
            with pytest.raises(_capi.PgtError):
                ctx.sync()
            ctx.upload(fixture_frames.position(199), big, box)
            ctx.build_cells()
            ctx.build_bonds()
            ctx.ik_compute(_capi.OBS_VIRIAL)
            got = ctx.ik_download()
            S = orc.System(box, fixture_frames.position(199), big)
            grid = np.zeros((64 * 64, 3)); grid[:, 0] = np.repeat(axis, 64); grid[:, 1] = np.tile(axis, 64)
            want = S.ik(P, orc.cgfunc("octic", 2.0, 5), grid, max_rcut=orc.System(box, fixture_frames.position(0), diam).max_rcut(P))["virial"]
            assert_close(got, want)
        ctx.close()
    for a, b in zip(fields[1], fields[0]):
        assert np.array_equal(a, b)


def test_two_contexts_on_one_device_keep_their_own_quadrature_rules(orc, fixture_frames):
    """Gauss-Legendre tables are per context (kernel arguments / per-context device buffers), never per device: two contexts
    with orders 5 and 7 -- and a third switched to the adaptive rule, which installs a 1-point table -- used in turn on
    one GPU each integrate with their own rule."""
    pos, diam, box = fixture_frames.position(150), fixture_frames.diameter, fixture_frames.box
    P = orc.potential("poly12")
    S = orc.System(box, pos, diam)
    L = float(box[0])
    axis = synth.linspace_axis(L, 40)
    grid = np.zeros((1600, 3)); grid[:, 0] = np.repeat(axis, 40); grid[:, 1] = np.tile(axis, 40)
    ctxs = {}
    for order in (5, 7):
        c = _capi.Context(0)
        c.set_potential("poly12", P.scaled_rcut, orc.pot_params(P))
        c.set_max_rcut(S.max_rcut(P))
        c.set_cg("octic", 2.0, order)
        c.set_grid_cartesian(axis, axis)
        ctxs[order] = c
    third = _capi.Context(0)
    third.set_potential("poly12", P.scaled_rcut, orc.pot_params(P))
    third.set_cg("octic", 2.0, 9)
    want = {order: S.ik(P, orc.cgfunc("octic", 2.0, order), grid)["virial"] for order in (5, 7)}
    for rep in range(2):
        for order in (7, 5):
            c = ctxs[order]
            c.upload(pos, diam, box); c.build_cells(); c.build_bonds()
            if rep == 1 and order == 7:
                third.set_cg_adaptive("octic", 2.0, 1e-9, 1e-9, 50)      # uploads a 1-point rule into ITS context only
            for variant in (0, 2):
                c.set_option("ik_variant", variant)
                c.ik_compute(_capi.OBS_VIRIAL)
                assert_close(c.ik_download(), want[order])
    assert np.abs(want[5] - want[7]).max() > 1e-9 * np.abs(want[5]).max()      # the two rules do differ
    for c in list(ctxs.values()) + [third]:
        c.close()


def test_symmetric_packed_readback(gpu_ctx, orc, fixture_frames):
    """pgt_ik_download_sym_async: (xx, xy, yy) per grid point equals columns 0, 1, 3 of the full read-back, and the
    dropped column 2 (yx) is bit-identical to column 1 in the full field."""
    P = orc.potential("poly12")
    pos, diam, box = fixture_frames.position(12), fixture_frames.diameter, fixture_frames.box
    setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    gpu_ctx.set_cg("octic", 2.0, 5)
    axis = synth.linspace_axis(float(box[0]), 50)
    gpu_ctx.set_grid_cartesian(axis, axis)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    full = gpu_ctx.ik_download()
    out = _capi.PinnedArray((2500, 3), np.float64)
    gpu_ctx.ik_download_sym_async(_capi.FIELD_VIRIAL, out.array)
    gpu_ctx.ik_download_wait()
    assert np.array_equal(out.array, full[:, [0, 1, 3]])
    assert np.array_equal(full[:, 1], full[:, 2])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_upload_xy_rows_give_the_same_bits_as_xyz_rows(gpu_ctx, dtype):
    """pgt_frame_upload_xy (N x 2 position rows of a 2-D frame) against pgt_frame_upload of the same frame with a zero z
    column: same live particles, same records, same field, bit for bit."""
    fr = synth.synth_frame(96, 2024, velocities=True)
    pos = fr["position"].astype(dtype)
    pos[7, 0] += dtype(fr["box"][0])                       # one particle outside the box: wrapped
    diam, vel = fr["diameter"].astype(dtype), fr["velocity"].astype(dtype)
    rc = 1.25
    gpu_ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    gpu_ctx.set_max_rcut(rc * float(diam.max()))
    gpu_ctx.set_cg("octic", 2.0, 5)
    axis = synth.linspace_axis(float(fr["box"][0]), 80)
    gpu_ctx.set_grid_cartesian(axis, axis)
    got = []
    for xy in (False, True):
        if xy:
            gpu_ctx.upload_xy(np.ascontiguousarray(pos[:, :2]), diam, fr["box"], vel=vel)
        else:
            gpu_ctx.upload(pos, diam, fr["box"], vel=vel)
        gpu_ctx.build_cells()
        gpu_ctx.build_bonds()
        gpu_ctx.ik_compute(_capi.OBS_VIRIAL | _capi.OBS_KINETIC)
        got.append((gpu_ctx.counts(), gpu_ctx.export_bonds(), gpu_ctx.ik_download(_capi.FIELD_VIRIAL).copy(),
                    gpu_ctx.ik_download(_capi.FIELD_KINETIC).copy(), gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_KINETIC)))
    a, b = got
    assert a[0] == b[0] and a[0]["n_live"] == len(diam)
    for u, v in zip(a[1], b[1]):
        assert np.array_equal(u, v)
    assert np.array_equal(a[2], b[2]) and np.abs(a[2]).max() > 0
    assert np.array_equal(a[3], b[3]) and np.abs(a[3]).max() > 0
    for k in ("virial", "kinetic"):
        assert np.array_equal(a[4][k], b[4][k])
    assert a[4]["energy"] == b[4]["energy"]


def test_crowded_cell_sorts_in_n_log_n_and_matches_the_oracle(gpu_ctx, orc):
    """A clustered frame: 1 024 particles inside one unit patch (a handful of cells hold hundreds of members each, the per-cell
    sort takes its heapsort path, every particle has far more than 16 bonds so the bond list takes its two-search path);
    live count, neighbour and bond counts, thermo sums and the field on a few grid points against the oracle."""
    rng = np.random.default_rng(31)
    L = 24.0
    bg = synth.synth_frame(24, 5)
    ii, jj = np.meshgrid(np.arange(32), np.arange(32), indexing="ij")
    patch = np.zeros((1024, 3), np.float32)
    patch[:, 0] = (3.0 + (ii.ravel() + 0.5) / 32 + rng.uniform(-0.004, 0.004, 1024)).astype(np.float32)
    patch[:, 1] = (-2.0 + (jj.ravel() + 0.5) / 32 + rng.uniform(-0.004, 0.004, 1024)).astype(np.float32)
    pos = np.concatenate([bg["position"], patch])
    order = rng.permutation(len(pos))                       # ids of the patch are scattered over the frame
    pos = pos[order]
    diam = np.concatenate([bg["diameter"], np.full(1024, 1.0, np.float32)])[order]
    box = np.array([L, L, 1.0])
    P = orc.potential("poly12")
    S = setup_ctx(gpu_ctx, orc, P, "poly12", pos, diam, box)
    gpu_ctx.build_bonds()
    cnt = gpu_ctx.counts()
    assert cnt["n_live"] == len(diam)
    oi, oj, _, ofl = S.neighbor_pairs(P)
    gi, gj = gpu_ctx.export_neighbors()
    assert len(gi) == len(oi) and pairs_key(gi, gj) == pairs_key(oi, oj)
    assert cnt["n_halfbonds"] == int(ofl.sum()) and cnt["n_halfbonds"] > 200 * 1024
    t = gpu_ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY)
    w = S.thermo(P, born=False)
    assert_close(t["virial"], w["virial"], atol=1e-12 * np.abs(w["virial"]).max())
    assert_close(t["energy"], w["energy"])
    gpu_ctx.set_cg("octic", 2.0, 5)
    grid = np.array([[3.5, -1.5, 0.0], [2.0, -3.0, 0.0], [-8.0, 7.0, 0.0], [5.2, -1.0, 0.0]])
    gpu_ctx.set_grid_points(grid)
    gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
    want = S.ik(P, orc.cgfunc("octic", 2.0, 5), grid)["virial"]
    assert_close(gpu_ctx.ik_download(_capi.FIELD_VIRIAL), want, atol=1e-12 * np.abs(want).max())


def test_thermo_async_over_a_trajectory_equals_the_synchronous_sums(orc):
    """pgt_thermo_compute_async in a loop that never waits (deferred counts): every frame's 25 doubles equal
    pgt_thermo_compute's bit for bit, and the first frame's equal the oracle's."""
    ctx = _capi.Context(0)
    try:
        frames = [synth.synth_frame(96, 300 + f, velocities=True) for f in range(6)]
        P = orc.potential("poly12")
        ctx.set_potential("poly12", P.scaled_rcut, orc.pot_params(P))
        dmax = max(float(f["diameter"].max()) for f in frames)
        ctx.set_max_rcut(max(P.scaled_rcut * dmax, dmax))
        mask = _capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_KINETIC | _capi.OBS_BORN
        want = []
        for fr in frames:
            ctx.upload(fr["position"], fr["diameter"], fr["box"], vel=fr["velocity"])
            ctx.build_cells()
            t = ctx.thermo(mask)
            want.append(np.concatenate([t["virial"], t["kinetic"], [t["energy"]], t["born"]]))
        outs = [_capi.PinnedArray((25,), np.float64) for _ in frames]
        for o in outs:
            o.array[...] = -1.0
        for rep in range(2):                                # the second pass runs with the counts deferred
            for fr, o in zip(frames, outs):
                ctx.upload(fr["position"], fr["diameter"], fr["box"], vel=fr["velocity"])
                ctx.build_cells()
                ctx.thermo_async(mask, o.array)
            ctx.sync()
            for o, w in zip(outs, want):
                assert np.array_equal(o.array, w)
        S = orc.System(frames[0]["box"], frames[0]["position"], frames[0]["diameter"], vel=frames[0]["velocity"])
        w0 = S.thermo(P)
        assert_close(want[0][0:4], w0["virial"])
        assert_close(want[0][8], w0["energy"])
    finally:
        ctx.close()


def test_host_register_locks_an_existing_array_in_place(gpu_ctx):
    """pgt_host_register / pgt_host_unregister: a frame uploaded from, and a field read back into, numpy arrays page-locked
    in place give the same bits as pageable ones; registering twice is refused, unregistering restores the array."""
    fr = synth.synth_frame(128, 99)
    rc = 1.25
    gpu_ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    gpu_ctx.set_max_rcut(rc * float(fr["diameter"].max()))
    gpu_ctx.set_cg("octic", 2.0, 5)
    axis = synth.linspace_axis(float(fr["box"][0]), 256)
    gpu_ctx.set_grid_cartesian(axis, axis)

    def field(pos, diam, out):
        gpu_ctx.upload(pos, diam, fr["box"])
        gpu_ctx.build_cells(); gpu_ctx.build_bonds(); gpu_ctx.ik_compute(_capi.OBS_VIRIAL)
        return gpu_ctx.ik_download(_capi.FIELD_VIRIAL, out)
    want = field(fr["position"], fr["diameter"], np.empty((256 * 256, 4))).copy()
    pos, diam, out = fr["position"].copy(), fr["diameter"].copy(), np.empty((256 * 256, 4))
    assert _capi.host_register(pos) and _capi.host_register(diam) and _capi.host_register(out)
    assert not _capi.host_register(out)                       # already registered
    try:
        got = field(pos, diam, out)
        assert np.array_equal(got, want)
    finally:
        assert _capi.host_unregister(pos) and _capi.host_unregister(diam) and _capi.host_unregister(out)
    assert not _capi.host_unregister(out)
    pos[0, 0] += 0.0                                          # the arrays are ordinary memory again
    assert np.array_equal(field(pos, diam, out), want)
