"""Pins the CPU oracle to the reference's own arithmetic (runs on CPU, no GPU needed).

* tests/golden/refmath_vectors.npz holds outputs of the REFERENCE'S OWN headers (potential/*.h,
  irvingkirkwood/cgfunc/CGFunc*.h compiled from /root/reference into oracle/_ref/libpgt_refmath.so by
  tests/golden/make_refmath_vectors.py).  The oracle must reproduce them bit for bit.
* When oracle/_ref is present (build container, or shipped prebuilt to the GPU box) the same
  comparison is repeated live on fresh random inputs.
* The Gauss-Legendre tables (GSL is not vendored: "parity unpinned" for the GSL call itself) are
  cross-checked against numpy's leggauss and against the defining properties of the rule.
"""
import os

import numpy as np
import pytest

MODELS = ["lj", "fslj", "poly12", "poly18", "poly10", "poly106", "polyyukawa", "polylj"]
CG = [("octic", 2.0), ("octic", 4.0), ("rect", 2.0), ("rect", 1.5), ("mollifier", 2.0), ("mollifier", 3.0)]
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "refmath_vectors.npz")


@pytest.fixture(scope="module")
def vectors():
    return np.load(GOLDEN)


@pytest.mark.parametrize("name", MODELS)
def test_pair_models_match_reference_headers_bitwise(orc, vectors, name):
    P = orc.potential(name)
    t = orc.pair_table(P, vectors["pair_%s_rij" % name], vectors["pair_%s_di" % name], vectors["pair_%s_dj" % name])
    for k in ("energy", "fdivr", "stiff", "rcutsq"):
        want = vectors["pair_%s_%s" % (name, k)]
        assert np.array_equal(t[k], want), "%s %s differs from the reference header output" % (name, k)
    # the table really straddles the cut-off: both zero and non-zero forces are present
    f = vectors["pair_%s_fdivr" % name]
    assert (f == 0).any() and (f != 0).any()


@pytest.mark.parametrize("kernel,R", CG)
def test_cg_kernels_match_reference_headers_bitwise(orc, vectors, kernel, R):
    cg = orc.cgfunc(kernel, R, 5)
    d2 = vectors["cg_%s_R%g_d2" % (kernel, R)]
    want = vectors["cg_%s_R%g_delta" % (kernel, R)]
    got = orc.cg_table(cg, d2)
    assert np.array_equal(got, want)
    assert (want == 0).any() and (want != 0).any()


@pytest.mark.parametrize("name", MODELS)
def test_pair_models_live_against_ref_lib(orc, name):
    if orc.ref_lib() is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine and no prebuilt copy)")
    rng = np.random.default_rng(hash(name) % 2**31)
    P = orc.potential(name)
    n = 4000
    di, dj = rng.uniform(0.7, 1.7, n), rng.uniform(0.7, 1.7, n)
    r = rng.uniform(0.4, 1.2, n) * P.scaled_rcut * 0.5 * (di + dj)
    th = rng.uniform(0, 2 * np.pi, n)
    rij = np.stack([r * np.cos(th), r * np.sin(th), np.zeros(n)], axis=1)
    a, b = orc.pair_table(P, rij, di, dj), orc.ref_pair_table(P, rij, di, dj)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 8, 12, 20, 32, 64])
def test_gauss_legendre_tables(orc, n):
    """GSL's glfixed layout: m=(n+1)/2 non-negative abscissae ascending, x0=0 for odd n."""
    x, w = orc.gl_table(n)
    xr, wr = np.polynomial.legendre.leggauss(n)
    m = (n + 1) // 2
    assert len(x) == m
    assert np.all(np.diff(x) > 0) and x[0] >= 0 and (n % 2 == 0 or x[0] == 0.0)
    np.testing.assert_allclose(x, xr[n - m:], rtol=0, atol=4e-16)
    np.testing.assert_allclose(w, wr[n - m:], rtol=0, atol=5e-15)    # numpy leggauss is itself only that accurate
    full_w = np.concatenate([w[::-1], w[1:] if n % 2 else w])
    assert abs(full_w.sum() - 2.0) < 1e-14
    # exact for polynomials up to degree 2n-1
    full_x = np.concatenate([-x[::-1], x[1:] if n % 2 else x])
    for deg in range(0, min(2 * n, 12), 2):
        assert abs((full_w * full_x**deg).sum() - 2.0 / (deg + 1)) < 1e-13


def test_gl5_hard_coded_values(orc):
    """The n=5 constants quoted in SURVEY.md section 8 (a8)."""
    x, w = orc.gl_table(5)
    np.testing.assert_allclose(x, [0.0, 0.5384693101056831, 0.9061798459386640], rtol=0, atol=1e-16)
    np.testing.assert_allclose(w, [0.5688888888888889, 0.4786286704993665, 0.2369268850561891], rtol=0, atol=1e-16)


@pytest.mark.parametrize("kernel", ["octic", "rect", "mollifier"])
@pytest.mark.parametrize("order", [1, 2, 5, 8, 20])
def test_bond_function_is_the_glfixed_rule(orc, kernel, order):
    """orc.bond == A * (w0 f(B) + sum w_i (f(B+A x_i) + f(B-A x_i))) on [0,1] with the oracle's own Delta
    (FixedPointQuadrature.h:35-45; GSL glfixed summation order)."""
    rng = np.random.default_rng(order)
    cg = orc.cgfunc(kernel, 2.0, order)
    x, w = orc.gl_table(order)
    for _ in range(50):
        gp = np.array([rng.uniform(-1, 1), rng.uniform(-1, 1), 0.0])
        ri = gp + np.array([rng.uniform(-2.2, 2.2), rng.uniform(-2.2, 2.2), 0.0])
        rij = np.array([rng.uniform(-1.5, 1.5), rng.uniform(-1.5, 1.5), 0.0])
        f = lambda s: orc.delta(cg, gp, ri + s * rij)
        A, B = 0.5, 0.5
        if order % 2:
            s = w[0] * f(B)
            rng_i = range(1, len(x))
        else:
            s = 0.0
            rng_i = range(len(x))
        for i in rng_i:
            Ax = A * x[i]
            s += w[i] * (f(B + Ax) + f(B - Ax))
        assert orc.bond(cg, gp, ri, rij) == A * s
