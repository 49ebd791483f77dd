// pgt_oracle.cpp -- CPU ORACLE for the IK / thermo pair-sum hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under pyglasstools_b200/ may include, link,
// dlopen or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, as the checker.
//
// It restates, loop for loop, the semantics of the reference (citations are
// into /root/reference/pyglasstools/):
//   * Aboria euclidean_search            extern/Aboria/src/Search.h:180-193,438-447
//   * domain enforcement                 extern/Aboria/src/NeighbourSearchBase.h:208-237
//   * SimBox bounds / volume             SimBox.cc:25-52
//   * Calculator::max_rcut               Calculator.h:37-46
//   * pair models                        potential/*.h, PairPotential.h:69-123
//   * CG kernels + bond integral         irvingkirkwood/cgfunc/CGFunc*.h, CoarseGrainFunction.h:56-92
//   * GSL glfixed rule (un-vendored GSL >= 2.0; call sites FixedPointQuadrature.h:35,45)
//   * IrvingKirkwood::compute            irvingkirkwood/IrvingKirkwood.h:74-110
//   * ThermoCalculator::compute          thermo/ThermoCalculator.h:68-90
//   * functors                           ListofCommonObservables.h:16-28,49-82,131-197
//
// Parity pin: the pair and CG arithmetic is checked bit-for-bit against the
// reference's own headers compiled into oracle/_ref/libpgt_refmath.so
// (tests/test_oracle_refmath.py, fixtures in tests/golden/), and the thermo path
// against the repository's only known-answer value (README.md:168-172).  The GSL
// call itself is "parity unpinned" (GSL is not vendored and absent here); its
// published rule is restated in orc_glfixed().
//
// Floating-point convention: this file is compiled with -ffp-contract=off, i.e.
// the ISO-C meaning of the reference's source (no FMA contraction).
#include <cmath>
#include <cfloat>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "gl_tables.h"

#define ORC_API extern "C"

// ---------------------------------------------------------------- pair models
// model ids (module-potential.cpp:13-20)
enum { ORC_LJ = 0, ORC_FSLJ = 1, ORC_POLY12 = 2, ORC_POLY18 = 3, ORC_POLY10 = 4,
       ORC_POLY106 = 5, ORC_POLYYUKAWA = 6, ORC_POLYLJ = 7 };

struct OrcPot { int model; double scaled_rcut; double p[8]; };

// what: 0 energy, 1 force/r, 2 stiffness, 3 rcut^2*sigma^2 (getRcut)
static double orc_model_eval(const OrcPot& P, double rsq, double d_i, double d_j, int what)
{
    const double rcutsq = P.scaled_rcut * P.scaled_rcut;       // PairPotential.h:101
    const double* q = P.p;
    switch (P.model) {
    case ORC_LJ:
    case ORC_FSLJ: {                                          // LennardJones.h
        const double eps = q[0], vareps = q[1];
        if (P.model == ORC_FSLJ && what == 1) {               // LennardJones.h:105-124 (name hiding)
            double sigma = 0.5 * (d_i + d_j);
            double actualrcutsq = rcutsq * sigma * sigma;
            if (rsq < actualrcutsq && eps != 0) {
                double r2inv = 1.0 / rsq;
                r2inv *= sigma * sigma;
                double r6inv = r2inv * r2inv * r2inv;
                double force_divr = 4 * eps * r2inv * r6inv * (12.0 * r6inv - 6.0);
                r2inv = 1.0 / rcutsq;
                r6inv = r2inv * r2inv * r2inv;
                force_divr -= 4 * eps * r2inv * r6inv * (12.0 * r6inv - 6.0);
                return force_divr / (sigma * sigma);
            }
            return 0.0;
        }
        double sigma = 0.5 * (d_i + d_j) * (1 - vareps * fabs(d_i - d_j));
        double actualrcutsq = rcutsq * sigma * sigma;
        if (what == 3) return actualrcutsq;
        if (what == 0) {                                      // LennardJones.h:19-38
            if (rsq < actualrcutsq) {
                double r2inv = 1.0 / rsq;
                r2inv *= sigma * sigma;
                double r6inv = r2inv * r2inv * r2inv;
                double r2invcut = 1.0 / actualrcutsq;
                r2invcut *= sigma * sigma;
                double r6invcut = r2invcut * r2invcut * r2invcut;
                return 4 * eps * r6inv * (r6inv - 1.0) - 4 * eps * r6invcut * (r6invcut - 1.0);
            }
            return 0.0;
        }
        if (rsq < actualrcutsq && eps != 0) {                 // :40-55, :57-72
            double r2inv = 1.0 / rsq;
            r2inv *= sigma * sigma;
            double r6inv = r2inv * r2inv * r2inv;
            if (what == 1) return (4 * eps * r2inv * r6inv * (12.0 * r6inv - 6.0)) / (sigma * sigma);
            return (4 * eps * r2inv * r6inv * (13.0 * 12.0 * r6inv - 6.0 * 7.0)) / (sigma * sigma);
        }
        return 0.0;
    }
    case ORC_POLYYUKAWA: {                                    // PolydisperseYukawa.h
        const double v0 = q[0], kappa = q[1], eps = q[2], c0 = q[3], c1 = q[4], c2 = q[5];
        double sigma = 0.5 * (d_i + d_j) * (1 - eps * fabs(d_i - d_j));
        double actualrcutsq = rcutsq * sigma * sigma;
        if (what == 3) return actualrcutsq;
        if (rsq < actualrcutsq) {
            double r2inv = 1.0 / rsq;
            r2inv *= sigma * sigma;
            double _rsq = rsq / (sigma * sigma);
            double rinv = sqrt(r2inv);
            double rroot = sqrt(_rsq);
            if (what == 0) return v0 * exp(-rroot * kappa) * rinv + c0 + c1 * _rsq + c2 * _rsq * _rsq;
            if (what == 1) return ((exp(-kappa * rroot) * (rinv + kappa) * v0) * r2inv - 2 * c1 - 4 * c2 * _rsq) / (sigma * sigma);
            return (v0 * exp(-kappa * rroot) * (2 * r2inv + 2 * kappa * rinv + kappa * kappa) * rinv + 2 * c1 + 12 * c2 * _rsq) / (sigma * sigma);
        }
        return 0.0;
    }
    default: break;
    }
    // the power-law polydisperse family: params [v0, eps, c0, c1, c2 (, c3)]
    const double v0 = q[0], eps = q[1], c0 = q[2], c1 = q[3], c2 = q[4], c3 = q[5];
    double sigma = 0.5 * (d_i + d_j) * (1 - eps * fabs(d_i - d_j));
    double actualrcutsq = rcutsq * sigma * sigma;
    if (what == 3) return actualrcutsq;
    if (!(rsq < actualrcutsq)) return 0.0;
    double r2inv = 1.0 / rsq;
    r2inv *= sigma * sigma;
    double _rsq = rsq / (sigma * sigma);
    double r6inv = r2inv * r2inv * r2inv;
    switch (P.model) {
    case ORC_POLY12:                                          // Polydisperse12.h:19-73
        if (what == 0) return v0 * r6inv * r6inv + c0 + c1 * _rsq + c2 * _rsq * _rsq;
        if (what == 1) return (12.0 * v0 * r2inv * r6inv * r6inv - 2.0 * c1 - 4.0 * c2 * _rsq) / (sigma * sigma);
        return (156.0 * v0 * r2inv * r6inv * r6inv + 2.0 * c1 + 12.0 * c2 * _rsq) / (sigma * sigma);
    case ORC_POLY18:                                          // Polydisperse18.h:19-72
        if (what == 0) return v0 * r6inv * r6inv * r6inv + c0 + c1 * _rsq + c2 * _rsq * _rsq;
        if (what == 1) return (18.0 * v0 * r2inv * r6inv * r6inv * r6inv - 2.0 * c1 - 4.0 * c2 * _rsq) / (sigma * sigma);
        return (342.0 * v0 * r2inv * r6inv * r6inv * r6inv + 2.0 * c1 + 12.0 * c2 * _rsq) / (sigma * sigma);
    case ORC_POLYLJ:                                          // PolydisperseLJ.h:19-72 (stiffness uses 4*c2)
        if (what == 0) return v0 * r6inv * r6inv - v0 * r6inv + c0 + c1 * _rsq + c2 * _rsq * _rsq;
        if (what == 1) return (12.0 * v0 * r2inv * r6inv * r6inv - 6.0 * v0 * r2inv * r6inv - 2.0 * c1 - 4.0 * c2 * _rsq) / (sigma * sigma);
        return (156.0 * v0 * r2inv * r6inv * r6inv - 42.0 * v0 * r2inv * r6inv + 2.0 * c1 + 4.0 * c2 * _rsq) / (sigma * sigma);
    case ORC_POLY10: {                                        // Polydisperse10.h:19-73
        double r10inv = r2inv * r2inv * r2inv * r2inv * r2inv;
        if (what == 0) return v0 * r10inv + c0 + c1 * _rsq + c2 * _rsq * _rsq + c3 * _rsq * _rsq * _rsq;
        if (what == 1) return (10.0 * v0 * r2inv * r10inv - 2.0 * c1 - 4.0 * c2 * _rsq - 6.0 * c3 * _rsq * _rsq) / (sigma * sigma);
        return (110.0 * v0 * r6inv * r6inv + 2.0 * c1 + 12.0 * c2 * _rsq + 30.0 * c3 * _rsq * _rsq) / (sigma * sigma);
    }
    case ORC_POLY106: {                                       // Polydisperse106.h:19-72
        double r10inv = r2inv * r2inv * r2inv * r2inv * r2inv;
        if (what == 0) return v0 * (r10inv - r2inv * r2inv * r2inv) + c0 + c1 * _rsq + c2 * _rsq * _rsq;
        if (what == 1) return (v0 * (10.0 * r10inv * r2inv - 6.0 * r2inv * r2inv * r2inv * r2inv) - 2.0 * c1 - 4.0 * c2 * _rsq) / (sigma * sigma);
        return (v0 * (110.0 * r6inv * r6inv - 42.0 * r2inv * r2inv * r2inv * r2inv) + 2.0 * c1 + 12.0 * c2 * _rsq) / (sigma * sigma);
    }
    }
    return 0.0;
}

// Eigen Vector3d::dot as used by PairPotential.h:100 / CoarseGrainFunction.h:80.
static inline double orc_dot3(const double* a, const double* b)
{
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

// ShortRangePairPotential<M>::getPairEnergy / getPairForceDivR / getBondStiffness / getRcut
ORC_API double orc_pair_eval(const OrcPot* P, const double* rij, double di, double dj, int what)
{
    return orc_model_eval(*P, orc_dot3(rij, rij), di, dj, what);
}

ORC_API void orc_pair_table(const OrcPot* P, int64_t n, const double* rij, const double* di, const double* dj,
                    double* energy, double* fdivr, double* stiff, double* rcutsq)
{
    for (int64_t k = 0; k < n; ++k) {
        energy[k] = orc_pair_eval(P, rij + 3 * k, di[k], dj[k], 0);
        fdivr[k] = orc_pair_eval(P, rij + 3 * k, di[k], dj[k], 1);
        stiff[k] = orc_pair_eval(P, rij + 3 * k, di[k], dj[k], 2);
        rcutsq[k] = orc_pair_eval(P, rij + 3 * k, di[k], dj[k], 3);
    }
}

// ------------------------------------------------------------ CG kernels
enum { ORC_OCTIC = 0, ORC_RECT = 1, ORC_MOLLIFIER = 2 };
struct OrcCG { int kernel; double rcut; int order; double mollifier_norm;
               int adaptive; double epsabs, epsrel; int limit; };   // AdaptiveCGFunc (CoarseGrainFunction.h:94-134)

ORC_API void orc_cg_init(OrcCG* cg, int kernel, double rcut, int order)
{
    cg->kernel = kernel; cg->rcut = rcut; cg->order = order;
    cg->adaptive = 0; cg->epsabs = cg->epsrel = 0.0; cg->limit = 0;
    // CGFuncMollifier.h:12 -- fixed at construction, not refreshed by setRcut
    cg->mollifier_norm = 13.46842098743079272180855772032086697297031125109170011102 / (2.0 * M_PI * rcut * rcut);
}

ORC_API double orc_cg_compute(const OrcCG* cg, double dr_sq)
{
    const double rcut = cg->rcut;
    double rcutsq = rcut * rcut;
    switch (cg->kernel) {
    case ORC_OCTIC:                                           // CGFuncOctic.h:16-29
        if (dr_sq < rcutsq) {
            double r4 = dr_sq * dr_sq / (rcutsq * rcutsq);
            double r8 = r4 * r4;
            return 15.0 / (8 * M_PI * rcutsq) * (1 - 2 * r4 + r8);
        }
        return 0.0;
    case ORC_RECT:                                            // CGFuncRect.h:17-28
        if (dr_sq < rcutsq) return 1 / (M_PI * rcutsq);
        return 0.0;
    case ORC_MOLLIFIER:                                       // CGFuncMollifier.h:20-32
        if (dr_sq < rcutsq) {
            double r2 = dr_sq / rcutsq;
            return cg->mollifier_norm * exp(-1.0 / (1 - r2));
        }
        return 0.0;
    }
    return 0.0;
}

// FixedPointCGFunc::getDeltaFunc (CoarseGrainFunction.h:77-82)
ORC_API double orc_delta(const OrcCG* cg, const double* x, const double* ri)
{
    double dr[3] = { x[0] - ri[0], x[1] - ri[1], x[2] - ri[2] };
    return orc_cg_compute(cg, orc_dot3(dr, dr));
}

// GSL gauss_legendre_tbl() for orders GSL does not tabulate (glfixed.c, eps = 1e-10).
static void orc_gl_newton(int n, double* x, double* w)
{
    const double eps = 1e-10;
    int m = (n + 1) >> 1;
    double t0 = (1.0 - (1.0 - 1.0 / (double)n) / (8.0 * (double)n * (double)n));
    double t1 = 1.0 / (4.0 * (double)n + 2.0);
    for (int i = 1; i <= m; i++) {
        double x0 = cos(3.1415926535897932384626433832795028841971693993751 * (double)((i << 2) - 1) * t1) * t0;
        double x1 = x0, w0 = 0, w1 = 0, dx, dw;
        int j = 0;
        do {
            double P_1 = 1.0, P0 = x0, P_2;
            for (int k = 2; k <= n; k++) {
                P_2 = P_1; P_1 = P0;
                double t2 = x0 * P_1;
                P0 = t2 + (1.0 - 1.0 / (double)k) * (t2 - P_2);
            }
            double dpdx = ((x0 * P0 - P_1) * (double)n) / (x0 * x0 - 1.0);
            x1 = x0 - P0 / dpdx;
            w1 = 2.0 / ((1.0 - x1 * x1) * dpdx * dpdx);
            if (j == 0) w0 = 2.0 / ((1.0 - x0 * x0) * dpdx * dpdx);
            dx = x0 - x1; dw = w0 - w1;
            x0 = x1; w0 = w1; j++;
        } while ((fabs(dx) > eps || fabs(dw) > eps) && j < 100);
        x[(m - 1) - (i - 1)] = x1;
        w[(m - 1) - (i - 1)] = w1;
    }
}

// gsl_integration_glfixed_table_alloc: returns m and fills x,w (capacity >= (n+1)/2)
ORC_API int orc_gl_table(int n, double* x, double* w)
{
    int m = (n + 1) >> 1;
    for (int t = 0; t < orc_gl_ntables; ++t)
        if (orc_gl_tables[t].n == n) {
            memcpy(x, orc_gl_tables[t].x, sizeof(double) * m);
            memcpy(w, orc_gl_tables[t].w, sizeof(double) * m);
            return m;
        }
    orc_gl_newton(n, x, w);
    return m;
}

struct OrcGL { int n, m; std::vector<double> x, w; };
static OrcGL orc_make_gl(int n)
{
    OrcGL g; g.n = n; g.m = (n + 1) >> 1; g.x.resize(g.m); g.w.resize(g.m);
    orc_gl_table(n, g.x.data(), g.w.data());
    return g;
}

// gsl_integration_glfixed(f, a, b, t)
template <class F>
static double orc_glfixed(F f, double a, double b, const OrcGL& t)
{
    const int n = t.n, m = t.m;
    const double A = 0.5 * (b - a), B = 0.5 * (b + a);
    double s;
    if (n & 1) {
        s = t.w[0] * f(B);
        for (int i = 1; i < m; i++) {
            double Ax = A * t.x[i];
            s += t.w[i] * (f(B + Ax) + f(B - Ax));
        }
    } else {
        s = 0.0;
        for (int i = 0; i < m; i++) {
            double Ax = A * t.x[i];
            s += t.w[i] * (f(B + Ax) + f(B - Ax));
        }
    }
    return A * s;
}

// FixedPointCGFunc::getBondFunc (CoarseGrainFunction.h:85-88) with getObjFunc (:42-45)
static double orc_bond_gl(const OrcCG* cg, const OrcGL& gl, const double* x, const double* ri, const double* rij)
{
    return orc_glfixed([&](double s) {
        double p[3] = { ri[0] + s * rij[0], ri[1] + s * rij[1], ri[2] + s * rij[2] };
        return orc_delta(cg, x, p);
    }, 0.0, 1.0, gl);
}

// ------------------------------------------------------------ adaptive quadrature (GSL QAG, key = 1)
// AdaptiveCGFunc::getBondFunc (CoarseGrainFunction.h:123-126) -> GSLQuadrature (Quadrature.h:39-51) ->
// gsl_integration_qag(f, 0, 1, epsabs, epsrel, limit, key = 1 (15-point Gauss-Kronrod), workspace).  GSL is an
// un-vendored dependency (>= 2.0, README.md:39): its published algorithm (integration/qag.c, qk.c, qk15.c, util.c,
// workspace.c; QUADPACK's DQAGE / DQK15) is restated here.  PARITY UNPINNED: the reference holds no value computed at
// this boundary; the restatement is checked against closed-form integrals and a high-order Gauss-Legendre rule.
ORC_API void orc_cg_init_adaptive(OrcCG* cg, int kernel, double rcut, double epsrel, double epsabs, int limit)
{
    orc_cg_init(cg, kernel, rcut, 0);
    cg->adaptive = 1; cg->epsabs = epsabs; cg->epsrel = epsrel; cg->limit = limit;
}

static const double orc_xgk15[8] = { 0.991455371120812639206854697526329, 0.949107912342758524526189684047851,
                                     0.864864423359769072789712788640926, 0.741531185599394439863864773280788,
                                     0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
                                     0.207784955007898467600689403773245, 0.000000000000000000000000000000000 };
static const double orc_wg15[4] = { 0.129484966168869693270611432679082, 0.279705391489276667901467771423780,
                                    0.381830050505118944950369775488975, 0.417959183673469387755102040816327 };
static const double orc_wgk15[8] = { 0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
                                     0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
                                     0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
                                     0.204432940075298892414161999234649, 0.209482141084727828012999174891714 };

ORC_API void orc_gk15_tables(double* xgk, double* wg, double* wgk)
{
    memcpy(xgk, orc_xgk15, sizeof(orc_xgk15)); memcpy(wg, orc_wg15, sizeof(orc_wg15)); memcpy(wgk, orc_wgk15, sizeof(orc_wgk15));
}

static double orc_rescale_error(double err, double result_abs, double result_asc)      // integration/err.c
{
    err = fabs(err);
    if (result_asc != 0 && err != 0) {
        double scale = pow((200 * err / result_asc), 1.5);
        err = (scale < 1) ? result_asc * scale : result_asc;
    }
    if (result_abs > DBL_MIN / (50 * DBL_EPSILON)) {
        double min_err = 50 * DBL_EPSILON * result_abs;
        if (min_err > err) err = min_err;
    }
    return err;
}

template <class F>
static void orc_qk15(F f, double a, double b, double* result, double* abserr, double* resabs, double* resasc)   // qk.c, n = 8
{
    const int n = 8;
    double fv1[8], fv2[8];
    const double center = 0.5 * (a + b), half_length = 0.5 * (b - a), abs_half_length = fabs(half_length);
    const double f_center = f(center);
    double result_gauss = 0, result_kronrod = f_center * orc_wgk15[n - 1], result_abs = fabs(result_kronrod), result_asc = 0;
    if (n % 2 == 0) result_gauss = f_center * orc_wg15[n / 2 - 1];
    for (int j = 0; j < (n - 1) / 2; j++) {
        const int jtw = j * 2 + 1;
        const double abscissa = half_length * orc_xgk15[jtw];
        const double fval1 = f(center - abscissa), fval2 = f(center + abscissa), fsum = fval1 + fval2;
        fv1[jtw] = fval1; fv2[jtw] = fval2;
        result_gauss += orc_wg15[j] * fsum;
        result_kronrod += orc_wgk15[jtw] * fsum;
        result_abs += orc_wgk15[jtw] * (fabs(fval1) + fabs(fval2));
    }
    for (int j = 0; j < n / 2; j++) {
        const int jtwm1 = j * 2;
        const double abscissa = half_length * orc_xgk15[jtwm1];
        const double fval1 = f(center - abscissa), fval2 = f(center + abscissa);
        fv1[jtwm1] = fval1; fv2[jtwm1] = fval2;
        result_kronrod += orc_wgk15[jtwm1] * (fval1 + fval2);
        result_abs += orc_wgk15[jtwm1] * (fabs(fval1) + fabs(fval2));
    }
    const double mean = result_kronrod * 0.5;
    result_asc = orc_wgk15[n - 1] * fabs(f_center - mean);
    for (int j = 0; j < n - 1; j++) result_asc += orc_wgk15[j] * (fabs(fv1[j] - mean) + fabs(fv2[j] - mean));
    const double err = (result_kronrod - result_gauss) * half_length;
    result_kronrod *= half_length; result_abs *= abs_half_length; result_asc *= abs_half_length;
    *result = result_kronrod; *resabs = result_abs; *resasc = result_asc;
    *abserr = orc_rescale_error(err, result_abs, result_asc);
}

// status: 0 converged; 1 round-off on the first interval; 2 limit reached; 3 round-off while bisecting; 4 interval too
// small.  (In the reference every non-zero status goes to GSL's default error handler, which aborts the process.)
template <class F>
static double orc_qag15(F f, double a, double b, double epsabs, double epsrel, int limit, int* status, double* abserr_out)
{
    std::vector<double> alist(1), blist(1), rlist(1), elist(1);
    double result0, abserr0, resabs0, resasc0;
    *status = 0;
    orc_qk15(f, a, b, &result0, &abserr0, &resabs0, &resasc0);
    alist[0] = a; blist[0] = b; rlist[0] = result0; elist[0] = abserr0;
    double tolerance = std::max(epsabs, epsrel * fabs(result0));
    const double round_off = 50 * DBL_EPSILON * resabs0;
    if (abserr_out) *abserr_out = abserr0;
    if (abserr0 <= round_off && abserr0 > tolerance) { *status = 1; return result0; }
    if ((abserr0 <= tolerance && abserr0 != resasc0) || abserr0 == 0.0) return result0;
    if (limit == 1) { *status = 2; return result0; }
    double area = result0, errsum = abserr0;
    int iteration = 1, roundoff_type1 = 0, roundoff_type2 = 0, error_type = 0;
    do {
        size_t imax = 0;                                    // workspace retrieve(): the interval with the largest error
        for (size_t k = 1; k < elist.size(); ++k) if (elist[k] > elist[imax]) imax = k;
        const double a_i = alist[imax], b_i = blist[imax], r_i = rlist[imax], e_i = elist[imax];
        const double a1 = a_i, b1 = 0.5 * (a_i + b_i), a2 = b1, b2 = b_i;
        double area1, area2, error1, error2, resabs1, resabs2, resasc1, resasc2;
        orc_qk15(f, a1, b1, &area1, &error1, &resabs1, &resasc1);
        orc_qk15(f, a2, b2, &area2, &error2, &resabs2, &resasc2);
        const double area12 = area1 + area2, error12 = error1 + error2;
        errsum += (error12 - e_i);
        area += area12 - r_i;
        if (resasc1 != error1 && resasc2 != error2) {
            const double delta = r_i - area12;
            if (fabs(delta) <= 1.0e-5 * fabs(area12) && error12 >= 0.99 * e_i) roundoff_type1++;
            if (iteration >= 10 && error12 > e_i) roundoff_type2++;
        }
        tolerance = std::max(epsabs, epsrel * fabs(area));
        if (errsum > tolerance) {
            if (roundoff_type1 >= 6 || roundoff_type2 >= 20) error_type = 2;
            const double tmp = (1 + 100 * DBL_EPSILON) * (fabs(a2) + 1000 * DBL_MIN);
            if (fabs(a1) <= tmp && fabs(b2) <= tmp) error_type = 3;
        }
        // workspace update(): the half with the larger error stays in place, the other is appended
        if (error2 > error1) {
            alist[imax] = a2; rlist[imax] = area2; elist[imax] = error2;
            alist.push_back(a1); blist.push_back(b1); rlist.push_back(area1); elist.push_back(error1);
        } else {
            blist[imax] = b1; rlist[imax] = area1; elist[imax] = error1;
            alist.push_back(a2); blist.push_back(b2); rlist.push_back(area2); elist.push_back(error2);
        }
        iteration++;
    } while (iteration < limit && !error_type && errsum > tolerance);
    double result = 0;
    for (size_t k = 0; k < rlist.size(); ++k) result += rlist[k];       // sum_results()
    if (abserr_out) *abserr_out = errsum;
    if (errsum <= tolerance) *status = 0;
    else if (error_type == 2) *status = 3;
    else if (error_type == 3) *status = 4;
    else if (iteration == limit) *status = 2;
    else *status = 5;
    return result;
}

static double orc_bond_any(const OrcCG* cg, const OrcGL& gl, const double* x, const double* ri, const double* rij, int* status)
{
    if (!cg->adaptive) return orc_bond_gl(cg, gl, x, ri, rij);
    int st = 0;
    const double v = orc_qag15([&](double s) {
        double p[3] = { ri[0] + s * rij[0], ri[1] + s * rij[1], ri[2] + s * rij[2] };
        return orc_delta(cg, x, p);
    }, 0.0, 1.0, cg->epsabs, cg->epsrel, cg->limit, &st, nullptr);
    if (st && status) *status = st;
    return v;
}

ORC_API double orc_bond(const OrcCG* cg, const double* x, const double* ri, const double* rij)
{
    OrcGL gl = orc_make_gl(cg->adaptive ? 1 : cg->order);
    return orc_bond_any(cg, gl, x, ri, rij, nullptr);
}

// gsl_integration_qag(key = 1) for a C callback: what oracle/stubs/gsl/gsl_integration.h hands the reference's
// cgfunc/Quadrature.h when the reference's own loops are compiled against the stand-ins (oracle/ref_loops_shim.cpp)
ORC_API int orc_gsl_qag15(double (*f)(double, void*), void* params, double a, double b, double epsabs, double epsrel, int limit,
                          double* result, double* abserr)
{
    int status = 0;
    double err = 0.0;
    *result = orc_qag15([&](double t) { return f(t, params); }, a, b, epsabs, epsrel, limit, &status, &err);
    if (abserr) *abserr = err;
    return status;
}

// generic test hook: integral of one of a few closed-form integrands over [a, b] with the QAG restatement
ORC_API double orc_qag_test(int which, double a, double b, double epsabs, double epsrel, int limit, int* status, double* abserr)
{
    auto f = [&](double t) {
        switch (which) {
        case 0: return exp(t);
        case 1: return 1.0 / (1.0 + 25.0 * t * t);
        case 2: return sqrt(fabs(t));
        case 3: return (t < 0.3) ? 1.0 : 0.0;
        default: return cos(40.0 * t);
        }
    };
    return orc_qag15(f, a, b, epsabs, epsrel, limit, status, abserr);
}

// ------------------------------------------------------------ particle system
struct OrcSystem {
    int dim;
    double boxsize[3], lower[3], upper[3], vol;
    int periodic[3];
    int64_t n;                       // live particles
    std::vector<double> pos, vel, diam, mass;   // pos/vel: n*3
    std::vector<double> disp;                  // n*3, zero until set (ParticleSystem.cc:28, setDisplacement :60-67)
    std::vector<int64_t> id;         // original index of each live particle
    // uniform search grid over x,y (z is a single slab)
    int ncx, ncy; double cw[2];
    std::vector<int64_t> cell_start, cell_items;
};

// SimBox(Vector3d boxsize, origin=0, dim) as built by utils.simbox (utils.py:9-22, SimBox.cc:25-52)
// + ParticleSystem ctor (ParticleSystem.cc:3-30) + Aboria enforce_domain (NeighbourSearchBase.h:208-237).
ORC_API OrcSystem* orc_system_create(int dim, const double* box, int64_t n, const double* pos, const double* vel,
                             const double* diam, const double* mass, double cell_hint)
{
    OrcSystem* S = new OrcSystem();
    S->dim = dim;
    for (int d = 0; d < 3; ++d) { S->boxsize[d] = box[d]; S->periodic[d] = 1; }
    if (dim == 2) S->boxsize[2] = 1.0;                       // utils.py:15-16
    for (int d = 0; d < 3; ++d) {
        S->upper[d] = 0.5 * S->boxsize[d] - 0.0;
        S->lower[d] = (-1.0) * S->upper[d] - 2.0 * 0.0;
    }
    if (dim == 2) { S->periodic[2] = 0; S->vol = S->boxsize[0] * S->boxsize[1]; }
    else S->vol = S->boxsize[0] * S->boxsize[1] * S->boxsize[2];

    for (int64_t i = 0; i < n; ++i) {
        double r[3] = { pos[3 * i], pos[3 * i + 1], pos[3 * i + 2] };
        bool alive = true;
        for (int d = 0; d < 3; ++d) {
            const double low = S->lower[d], high = S->upper[d];
            if (!std::isfinite(r[d])) alive = false;
            else if (S->periodic[d]) {
                while (r[d] < low) r[d] += (high - low);
                while (r[d] >= high) r[d] -= (high - low);
            } else if ((r[d] < low) || (r[d] >= high)) alive = false;
        }
        if (!alive) continue;
        S->pos.insert(S->pos.end(), r, r + 3);
        for (int d = 0; d < 3; ++d) S->vel.push_back(vel ? vel[3 * i + d] : 0.0);
        S->diam.push_back(diam[i]);
        S->mass.push_back(mass ? mass[i] : 1.0);
        S->id.push_back(i);
    }
    S->n = (int64_t)S->id.size();
    // search grid (an implementation detail: enumeration order only changes summation order)
    double side = cell_hint > 0 ? cell_hint : 1.0;
    S->ncx = std::max(1, (int)floor(S->boxsize[0] / side));
    S->ncy = std::max(1, (int)floor(S->boxsize[1] / side));
    S->cw[0] = S->boxsize[0] / S->ncx; S->cw[1] = S->boxsize[1] / S->ncy;
    std::vector<int64_t> cnt((size_t)S->ncx * S->ncy + 1, 0);
    std::vector<int> cell(S->n);
    for (int64_t i = 0; i < S->n; ++i) {
        int cx = std::min(S->ncx - 1, std::max(0, (int)floor((S->pos[3 * i] - S->lower[0]) / S->cw[0])));
        int cy = std::min(S->ncy - 1, std::max(0, (int)floor((S->pos[3 * i + 1] - S->lower[1]) / S->cw[1])));
        cell[i] = cy * S->ncx + cx; cnt[cell[i] + 1]++;
    }
    for (size_t c = 1; c < cnt.size(); ++c) cnt[c] += cnt[c - 1];
    S->cell_start = cnt;
    S->cell_items.resize(S->n);
    std::vector<int64_t> fill(cnt.begin(), cnt.end() - 1);
    for (int64_t i = 0; i < S->n; ++i) S->cell_items[fill[cell[i]]++] = i;
    return S;
}
ORC_API void orc_system_destroy(OrcSystem* S) { delete S; }
ORC_API int64_t orc_system_n(const OrcSystem* S) { return S->n; }
ORC_API double orc_system_vol(const OrcSystem* S) { return S->vol; }
ORC_API void orc_system_get(const OrcSystem* S, double* pos, int64_t* id)
{
    if (pos) memcpy(pos, S->pos.data(), sizeof(double) * 3 * S->n);
    if (id) memcpy(id, S->id.data(), sizeof(int64_t) * S->n);
}

// Calculator ctor (Calculator.h:37-46)
ORC_API double orc_max_rcut(const OrcSystem* S, double scaled_rcut)
{
    double maxforce_rcut = scaled_rcut;
    double maxdiameter = *std::max_element(S->diam.begin(), S->diam.end());
    maxforce_rcut *= maxdiameter;
    return std::max(maxforce_rcut, maxdiameter);
}

// Aboria euclidean_search (Search.h:180-193 image loop, :438-447 candidate test,
// detail/Distance.h norm2: accum = ((0 + dx0^2) + dx1^2) + dx2^2, inclusive <=).
template <class F>
static inline void orc_search(const OrcSystem& S, const double* c, double radius, F&& hit)
{
    const double r2 = radius * radius;                        // std::pow(max_distance, 2)
    const int k0[3] = { S.periodic[0] ? -1 : 0, S.periodic[1] ? -1 : 0, S.periodic[2] ? -1 : 0 };
    const int k1[3] = { S.periodic[0] ? 2 : 1, S.periodic[1] ? 2 : 1, S.periodic[2] ? 2 : 1 };
    for (int kz = k0[2]; kz < k1[2]; ++kz)
    for (int ky = k0[1]; ky < k1[1]; ++ky)
    for (int kx = k0[0]; kx < k1[0]; ++kx) {
        const int k[3] = { kx, ky, kz };
        double cp[3];
        for (int d = 0; d < 3; ++d) cp[d] = c[d] + k[d] * (S.upper[d] - S.lower[d]);
        // buckets near the shifted point, clipped to the domain
        const double pad = 1e-9;
        int cx0 = (int)floor((cp[0] - radius - S.lower[0]) / S.cw[0] - pad);
        int cx1 = (int)floor((cp[0] + radius - S.lower[0]) / S.cw[0] + pad);
        int cy0 = (int)floor((cp[1] - radius - S.lower[1]) / S.cw[1] - pad);
        int cy1 = (int)floor((cp[1] + radius - S.lower[1]) / S.cw[1] + pad);
        cx0 = std::max(cx0, 0); cy0 = std::max(cy0, 0);
        cx1 = std::min(cx1, S.ncx - 1); cy1 = std::min(cy1, S.ncy - 1);
        for (int cy = cy0; cy <= cy1; ++cy)
        for (int cx = cx0; cx <= cx1; ++cx) {
            const int64_t cb = S.cell_start[(size_t)cy * S.ncx + cx], ce = S.cell_start[(size_t)cy * S.ncx + cx + 1];
            for (int64_t q = cb; q < ce; ++q) {
                const int64_t p = S.cell_items[q];
                double dx[3] = { S.pos[3 * p] - cp[0], S.pos[3 * p + 1] - cp[1], S.pos[3 * p + 2] - cp[2] };
                double accum = 0;
                accum = accum + dx[0] * dx[0];
                accum = accum + dx[1] * dx[1];
                accum = accum + dx[2] * dx[2];
                if (!(accum > r2)) hit(p, dx);
            }
        }
    }
}

// brute-force variant of the same search (no buckets) used to validate orc_search itself
template <class F>
static inline void orc_search_brute(const OrcSystem& S, const double* c, double radius, F&& hit)
{
    const double r2 = radius * radius;
    for (int kz = (S.periodic[2] ? -1 : 0); kz < (S.periodic[2] ? 2 : 1); ++kz)
    for (int ky = (S.periodic[1] ? -1 : 0); ky < (S.periodic[1] ? 2 : 1); ++ky)
    for (int kx = (S.periodic[0] ? -1 : 0); kx < (S.periodic[0] ? 2 : 1); ++kx) {
        const int k[3] = { kx, ky, kz };
        double cp[3];
        for (int d = 0; d < 3; ++d) cp[d] = c[d] + k[d] * (S.upper[d] - S.lower[d]);
        for (int64_t p = 0; p < S.n; ++p) {
            double dx[3] = { S.pos[3 * p] - cp[0], S.pos[3 * p + 1] - cp[1], S.pos[3 * p + 2] - cp[2] };
            double accum = 0;
            accum = accum + dx[0] * dx[0];
            accum = accum + dx[1] * dx[1];
            accum = accum + dx[2] * dx[2];
            if (!(accum > r2)) hit(p, dx);
        }
    }
}

// Neighbour set of the reference's inner search: ordered pairs (id_i, id_j), i != j,
// with ||r_j - (r_i + kL)|| <= max_rcut.  flags bit0: the pair carries force (rsq < rcut^2 sigma^2).
// Returns the count; writes up to cap entries.  (IrvingKirkwood.h:93-101 / ThermoCalculator.h:77-84)
ORC_API int64_t orc_neighbor_pairs(const OrcSystem* S, const OrcPot* P, double max_rcut, int brute,
                           int64_t cap, int64_t* pi, int64_t* pj, double* rij_out, int8_t* flags)
{
    int64_t cnt = 0;
    for (int64_t i = 0; i < S->n; ++i) {
        auto visit = [&](int64_t j, const double* dx) {
            if (S->id[i] == S->id[j]) return;
            double rij[3] = { -dx[0], -dx[1], -dx[2] };
            if (cnt < cap) {
                pi[cnt] = S->id[i]; pj[cnt] = S->id[j];
                if (rij_out) { rij_out[3 * cnt] = rij[0]; rij_out[3 * cnt + 1] = rij[1]; rij_out[3 * cnt + 2] = rij[2]; }
                if (flags) {
                    double rsq = orc_dot3(rij, rij);
                    flags[cnt] = (rsq < orc_model_eval(*P, rsq, S->diam[i], S->diam[j], 3)) ? 1 : 0;
                }
            }
            ++cnt;
        };
        if (brute) orc_search_brute(*S, &S->pos[3 * i], max_rcut, visit);
        else orc_search(*S, &S->pos[3 * i], max_rcut, visit);
    }
    return cnt;
}

// ------------------------------------------------------------ IrvingKirkwood::compute
// obs_mask: bit0 virial stress field, bit1 kinetic stress field, bit2 density (sum of Delta; not in the
// reference -- north_star stage 5).  kinetic_mode: 0 = intended  sum_i m v(x)v Delta  (what the GPU
// implements), 1 = the reference's order-dependent recurrence (ListofCommonObservables.h:143-148).
// Outputs (any may be NULL): virial[G*Dim*Dim] col-major per point (index a + Dim*b), kinetic likewise,
// density[G]; counts[G*3] = {particles in disc, searched triples, force-carrying ordered triples}.
ORC_API void orc_ik_compute(const OrcSystem* S, const OrcPot* P, const OrcCG* cg, double max_rcut,
                    int64_t G, const double* grid, unsigned obs_mask, int kinetic_mode, int brute,
                    double* virial, double* kinetic, double* density, int64_t* counts)
{
    const int Dim = S->dim;
    const OrcGL gl = orc_make_gl(cg->adaptive ? 1 : cg->order);
    const double cg_rcut = cg->rcut;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t g = 0; g < G; ++g) {
        const double* x = grid + 3 * g;
        double val[9] = { 0 }, kin[9] = { 0 }, rho = 0;       // clearState (:79)
        int64_t c_part = 0, c_trip = 0, c_force = 0;
        auto outer = [&](int64_t i, const double* dr) {          // :80-110
            double ri[3] = { x[0] - dr[0], x[1] - dr[1], x[2] - dr[2] };
            double cgval = orc_delta(cg, x, ri);               // :88
            ++c_part;
            rho += cgval;
            if (obs_mask & 2u) {
                const double m = S->mass[i]; const double* v = &S->vel[3 * i];
                if (kinetic_mode == 0) {
                    for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a)
                        kin[a + Dim * b] += (m * v[a] * v[b]) * cgval;
                } else {                                       // KineticStress::compute_cg
                    for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a)
                        kin[a + Dim * b] += m * v[a] * v[b];
                    for (int q = 0; q < Dim * Dim; ++q) kin[q] = kin[q] * cgval;
                }
            }
            if (!(obs_mask & 1u) && !counts) return;
            auto inner = [&](int64_t j, const double* dxj) {     // :93-107
                if (S->id[i] == S->id[j]) return;
                ++c_trip;
                double rij[3] = { -dxj[0], -dxj[1], -dxj[2] };
                double forceval = orc_model_eval(*P, orc_dot3(rij, rij), S->diam[i], S->diam[j], 1);
                if (forceval != 0.0) ++c_force;
                if (!(obs_mask & 1u)) return;
                double bondval = orc_bond_any(cg, gl, x, ri, rij, nullptr); // :102
                // VirialStress::compute_cg (ListofCommonObservables.h:72-82) on a zeroed temporary
                double F[3] = { forceval * rij[0], forceval * rij[1], forceval * rij[2] };
                const double half_bond = 0.5 * bondval;
                for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a) {
                    double tmp = 0.0; tmp += F[a] * rij[b];
                    val[a + Dim * b] += half_bond * tmp;
                }
            };
            if (brute) orc_search_brute(*S, &S->pos[3 * i], max_rcut, inner);
            else orc_search(*S, &S->pos[3 * i], max_rcut, inner);
        };
        if (brute) orc_search_brute(*S, x, cg_rcut, outer);
        else orc_search(*S, x, cg_rcut, outer);
        if (virial) for (int q = 0; q < Dim * Dim; ++q) virial[g * Dim * Dim + q] = val[q];
        if (kinetic) for (int q = 0; q < Dim * Dim; ++q) kinetic[g * Dim * Dim + q] = kin[q];
        if (density) density[g] = rho;
        if (counts) { counts[3 * g] = c_part; counts[3 * g + 1] = c_trip; counts[3 * g + 2] = c_force; }
    }
}

// U(x): unique force-carrying bonds {i,j} with at least one endpoint within cg_rcut of x
// (the algorithmic work unit of the IK gather).
ORC_API void orc_ik_unique_bonds(const OrcSystem* S, const OrcPot* P, const OrcCG* cg, double max_rcut,
                         int64_t G, const double* grid, int64_t* U)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t g = 0; g < G; ++g) {
        const double* x = grid + 3 * g;
        std::vector<int64_t> inside;
        orc_search(*S, x, cg->rcut, [&](int64_t i, const double*) { inside.push_back(i); });
        std::sort(inside.begin(), inside.end());
        int64_t u = 0;
        for (int64_t i : inside)
            orc_search(*S, &S->pos[3 * i], max_rcut, [&](int64_t j, const double* dxj) {
                if (i == j) return;
                double rij[3] = { -dxj[0], -dxj[1], -dxj[2] };
                if (orc_model_eval(*P, orc_dot3(rij, rij), S->diam[i], S->diam[j], 1) == 0.0) return;
                bool j_inside = std::binary_search(inside.begin(), inside.end(), j);
                if (!j_inside || i < j) ++u;
            });
        U[g] = u;
    }
}

// ------------------------------------------------------------ ThermoCalculator::compute
// out: virial[Dim*Dim], kinetic[Dim*Dim], born[Dim^4] (col-major flat, index i + Dim*(j + Dim*(k + Dim*l))),
// energy[1]; everything multiplied by (1/vol) (ThermoCalculator.h:58-64,89).  Serial, particle order.
ORC_API void orc_thermo_compute(const OrcSystem* S, const OrcPot* P, double max_rcut, int brute,
                        double* virial, double* kinetic, double* born, double* energy, int64_t* npairs)
{
    const int Dim = S->dim;
    double vir[9] = { 0 }, kin[9] = { 0 }, bt[81] = { 0 }, en = 0;
    int64_t np = 0;
    for (int64_t i = 0; i < S->n; ++i) {
        {   // KineticStress::compute (:131-142)
            const double m = S->mass[i]; const double* v = &S->vel[3 * i];
            for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a) kin[a + Dim * b] += m * v[a] * v[b];
        }
        auto inner = [&](int64_t j, const double* dx) {
            if (!(S->id[i] < S->id[j])) return;                 // :81
            ++np;
            const double rij[3] = { dx[0], dx[1], dx[2] };       // :84
            const double rsq = orc_dot3(rij, rij);
            const double di = S->diam[i], dj = S->diam[j];
            const double forceval = orc_model_eval(*P, rsq, di, dj, 1);
            const double F[3] = { forceval * rij[0], forceval * rij[1], forceval * rij[2] };
            for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a) vir[a + Dim * b] += F[a] * rij[b];
            en += orc_model_eval(*P, rsq, di, dj, 0);
            if (born) {                                        // BornStiffnessTensor::compute (:157-184)
                const double stiffnessval = orc_model_eval(*P, rsq, di, dj, 2);
                const double nrm = sqrt(rsq);
                const double nij[3] = { rij[0] / nrm, rij[1] / nrm, rij[2] / nrm };
                for (int l = 0; l < Dim; ++l) for (int k = 0; k < Dim; ++k)
                for (int j2 = 0; j2 < Dim; ++j2) for (int i2 = 0; i2 < Dim; ++i2)
                    bt[i2 + Dim * (j2 + Dim * (k + Dim * l))] += (stiffnessval * (rij[i2]) + F[i2]) * rij[j2] * nij[k] * nij[l];
            }
        };
        if (brute) orc_search_brute(*S, &S->pos[3 * i], max_rcut, inner);
        else orc_search(*S, &S->pos[3 * i], max_rcut, inner);
    }
    const double invvol = 1 / S->vol;                          // ThermoProperty.h:84-87
    if (virial) for (int q = 0; q < Dim * Dim; ++q) virial[q] = vir[q] * invvol;
    if (kinetic) for (int q = 0; q < Dim * Dim; ++q) kinetic[q] = kin[q] * invvol;
    if (born) for (int q = 0; q < Dim * Dim * Dim * Dim; ++q) born[q] = bt[q] * invvol;
    if (energy) energy[0] = en * invvol;
    if (npairs) npairs[0] = np;
}

// ------------------------------------------------------------ ElasticVirialStress (ListofCommonObservables.h:86-123)
// ParticleSystem::setDisplacement (ParticleSystem.cc:60-67): one row per ORIGINAL particle.
ORC_API void orc_system_set_displacement(OrcSystem* S, const double* disp)
{
    S->disp.assign((size_t)3 * S->n, 0.0);
    if (!disp) return;
    for (int64_t i = 0; i < S->n; ++i)
        for (int d = 0; d < 3; ++d) S->disp[3 * i + d] = disp[3 * S->id[i] + d];
}

// ElasticVirialStress::compute: val(a,b) += F1[a] (uij+rij)[b] - F[a] rij[b],  uij = disp_j - disp_i,
// F = fdivr(rij) rij, F1 = fdivr(uij + rij) (uij + rij)
static inline void orc_elastic_pair(const OrcSystem* S, const OrcPot* P, int64_t i, int64_t j, const double* rij, int Dim, double* tmp)
{
    const double* ui = S->disp.empty() ? nullptr : &S->disp[3 * i];
    const double* uj = S->disp.empty() ? nullptr : &S->disp[3 * j];
    double uij[3] = { 0, 0, 0 };
    if (ui) for (int d = 0; d < 3; ++d) uij[d] = uj[d] - ui[d];
    const double w[3] = { uij[0] + rij[0], uij[1] + rij[1], uij[2] + rij[2] };
    const double di = S->diam[i], dj = S->diam[j];
    const double forceval = orc_model_eval(*P, orc_dot3(rij, rij), di, dj, 1);
    const double forceval1 = orc_model_eval(*P, orc_dot3(w, w), di, dj, 1);
    const double F[3] = { forceval * rij[0], forceval * rij[1], forceval * rij[2] };
    const double F1[3] = { forceval1 * w[0], forceval1 * w[1], forceval1 * w[2] };
    for (int b = 0; b < Dim; ++b) for (int a = 0; a < Dim; ++a) tmp[a + Dim * b] += F1[a] * (uij[b] + rij[b]) - F[a] * rij[b];
}

// GlobalElasticVirialStress: ThermoCalculator::compute (ThermoCalculator.h:68-90), rij = +dx
ORC_API void orc_thermo_elastic(const OrcSystem* S, const OrcPot* P, double max_rcut, double* out)
{
    const int Dim = S->dim;
    double val[9] = { 0 };
    for (int64_t i = 0; i < S->n; ++i)
        orc_search(*S, &S->pos[3 * i], max_rcut, [&](int64_t j, const double* dx) {
            if (!(S->id[i] < S->id[j])) return;
            const double rij[3] = { dx[0], dx[1], dx[2] };
            orc_elastic_pair(S, P, i, j, rij, Dim, val);
        });
    const double invvol = 1 / S->vol;
    for (int q = 0; q < Dim * Dim; ++q) out[q] = val[q] * invvol;
}

// ElasticVirialStressField: IrvingKirkwood::compute (IrvingKirkwood.h:74-110) hands the functor rij = -dx (:100), so
// the displaced separation it evaluates is  (disp_j - disp_i) + (r_i - r_j)  -- reproduced as written.
ORC_API void orc_ik_elastic(const OrcSystem* S, const OrcPot* P, const OrcCG* cg, double max_rcut, int64_t G, const double* grid,
                            double* out)
{
    const int Dim = S->dim;
    const OrcGL gl = orc_make_gl(cg->adaptive ? 1 : cg->order);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t g = 0; g < G; ++g) {
        const double* x = grid + 3 * g;
        double val[9] = { 0 };
        orc_search(*S, x, cg->rcut, [&](int64_t i, const double* dr) {
            const double ri[3] = { x[0] - dr[0], x[1] - dr[1], x[2] - dr[2] };
            orc_search(*S, &S->pos[3 * i], max_rcut, [&](int64_t j, const double* dxj) {
                if (S->id[i] == S->id[j]) return;
                const double rij[3] = { -dxj[0], -dxj[1], -dxj[2] };
                const double bondval = orc_bond_any(cg, gl, x, ri, rij, nullptr);
                double tmp[9] = { 0 };                          // compute_cg: zeroed temporary, then 0.5 bondval tmp
                orc_elastic_pair(S, P, i, j, rij, Dim, tmp);
                for (int q = 0; q < Dim * Dim; ++q) val[q] += 0.5 * bondval * tmp[q];
            });
        });
        for (int q = 0; q < Dim * Dim; ++q) out[g * Dim * Dim + q] = val[q];
    }
}

ORC_API int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
ORC_API void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

