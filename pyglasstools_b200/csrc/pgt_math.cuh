// pgt_math.cuh -- pair models and coarse-graining shapes, shared by device kernels and the
// scalar host getters of the C-ABI.
//
// The whole library is compiled with -fmad=false (device) / -ffp-contract=off (host): a*b+c is
// never silently fused, so the discrete decisions (rsq < rcut^2 sigma^2, |dx|^2 <= R^2) are the
// ISO-C reading of the reference source and bit-identical between host, device and the oracle.
// Fused multiply-adds are written explicitly (fma()) where we want them (the IK quadrature).
//
// Reference semantics restated here (file:line under /root/reference/pyglasstools/):
//   potential/PairPotential.h:92-122, potential/{LennardJones,Polydisperse12,Polydisperse18,
//   PolydisperseLJ,Polydisperse10,Polydisperse106,PolydisperseYukawa}.h,
//   irvingkirkwood/cgfunc/CGFunc{Octic,Rect,Mollifier}.h
#pragma once
#include <math.h>
#include "pgt_cuda.h"

#if defined(__CUDACC__)
#define PGT_HD __host__ __device__ __forceinline__
#else
#define PGT_HD inline
#endif

#if defined(__CUDA_ARCH__)
#define PGT_UNROLL _Pragma("unroll")
#else
#define PGT_UNROLL
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

struct PgtPot {
    int model;
    int nparams;
    double scaled_rcut;
    double rcutsq;      // scaled_rcut * scaled_rcut
    double p[8];
};

enum { PGT_WANT_ENERGY = 0, PGT_WANT_FDIVR = 1, PGT_WANT_STIFF = 2, PGT_WANT_RCUTSQ = 3 };

// Additive-mixing diameter with non-additivity eps:  0.5 (di+dj) (1 - eps |di-dj|)
PGT_HD double pgt_sigma(double di, double dj, double eps)
{
    return 0.5 * (di + dj) * (1 - eps * fabs(di - dj));
}

// Index of the non-additivity parameter inside params for each model (the LJ pair keeps it at [1]).
PGT_HD int pgt_eps_index(int model) { return model == PGT_POT_POLYYUKAWA ? 2 : 1; }

// rcut^2 sigma^2 for the pair, i.e. the squared distance below which the pair carries force.
// (force-shifted LJ tests the force against the additive sigma, LennardJones.h:111-112.)
template <int MODEL>
PGT_HD double pgt_pair_cutsq(const PgtPot& P, double di, double dj, bool for_force)
{
    double sigma = (MODEL == PGT_POT_FORCESHIFTED_LJ && for_force) ? 0.5 * (di + dj)
                                                                 : pgt_sigma(di, dj, P.p[pgt_eps_index(MODEL)]);
    return P.rcutsq * sigma * sigma;
}

// True when the ordered pair (rsq, di, dj) contributes a non-zero force in the reference.
template <int MODEL>
PGT_HD bool pgt_pair_has_force(const PgtPot& P, double rsq, double di, double dj)
{
    if ((MODEL == PGT_POT_LJ || MODEL == PGT_POT_FORCESHIFTED_LJ) && P.p[0] == 0) return false;
    return rsq < pgt_pair_cutsq<MODEL>(P, di, dj, true);
}

// One evaluator for all eight models.  u = sigma^2/rsq ("r2inv"), t = rsq/sigma^2 ("_rsq").
template <int MODEL>
PGT_HD double pgt_pair_eval_t(const PgtPot& P, double rsq, double di, double dj, int what)
{
    if (what == PGT_WANT_RCUTSQ) return pgt_pair_cutsq<MODEL>(P, di, dj, false);

    if (MODEL == PGT_POT_LJ || MODEL == PGT_POT_FORCESHIFTED_LJ) {
        const double eps = P.p[0];
        if (MODEL == PGT_POT_FORCESHIFTED_LJ && what == PGT_WANT_FDIVR) {
            const double sigma = 0.5 * (di + dj);
            if (!(rsq < P.rcutsq * sigma * sigma && eps != 0)) return 0.0;
            double u = 1.0 / rsq;
            u *= sigma * sigma;
            double u3 = u * u * u;
            double f = 4 * eps * u * u3 * (12.0 * u3 - 6.0);
            u = 1.0 / P.rcutsq;
            u3 = u * u * u;
            f -= 4 * eps * u * u3 * (12.0 * u3 - 6.0);
            return f / (sigma * sigma);
        }
        const double sigma = pgt_sigma(di, dj, P.p[1]);
        const double cutsq = P.rcutsq * sigma * sigma;
        if (!(rsq < cutsq)) return 0.0;
        double u = 1.0 / rsq;
        u *= sigma * sigma;
        const double u3 = u * u * u;
        if (what == PGT_WANT_ENERGY) {
            double uc = 1.0 / cutsq;
            uc *= sigma * sigma;
            const double uc3 = uc * uc * uc;
            return 4 * eps * u3 * (u3 - 1.0) - 4 * eps * uc3 * (uc3 - 1.0);
        }
        if (eps == 0) return 0.0;
        if (what == PGT_WANT_FDIVR) return (4 * eps * u * u3 * (12.0 * u3 - 6.0)) / (sigma * sigma);
        return (4 * eps * u * u3 * (13.0 * 12.0 * u3 - 6.0 * 7.0)) / (sigma * sigma);
    }

    const double sigma = pgt_sigma(di, dj, P.p[pgt_eps_index(MODEL)]);
    if (!(rsq < P.rcutsq * sigma * sigma)) return 0.0;
    double u = 1.0 / rsq;
    u *= sigma * sigma;
    const double t = rsq / (sigma * sigma);
    const double ss = sigma * sigma;

    if (MODEL == PGT_POT_POLYYUKAWA) {
        const double v0 = P.p[0], kappa = P.p[1], c0 = P.p[3], c1 = P.p[4], c2 = P.p[5];
        const double rinv = sqrt(u), rroot = sqrt(t);
        if (what == PGT_WANT_ENERGY) return v0 * exp(-rroot * kappa) * rinv + c0 + c1 * t + c2 * t * t;
        if (what == PGT_WANT_FDIVR) return ((exp(-kappa * rroot) * (rinv + kappa) * v0) * u - 2 * c1 - 4 * c2 * t) / ss;
        return (v0 * exp(-kappa * rroot) * (2 * u + 2 * kappa * rinv + kappa * kappa) * rinv + 2 * c1 + 12 * c2 * t) / ss;
    }

    const double v0 = P.p[0], c0 = P.p[2], c1 = P.p[3], c2 = P.p[4];
    const double u3 = u * u * u;
    if (MODEL == PGT_POT_POLY12) {
        if (what == PGT_WANT_ENERGY) return v0 * u3 * u3 + c0 + c1 * t + c2 * t * t;
        if (what == PGT_WANT_FDIVR) return (12.0 * v0 * u * u3 * u3 - 2.0 * c1 - 4.0 * c2 * t) / ss;
        return (156.0 * v0 * u * u3 * u3 + 2.0 * c1 + 12.0 * c2 * t) / ss;
    }
    if (MODEL == PGT_POT_POLY18) {
        if (what == PGT_WANT_ENERGY) return v0 * u3 * u3 * u3 + c0 + c1 * t + c2 * t * t;
        if (what == PGT_WANT_FDIVR) return (18.0 * v0 * u * u3 * u3 * u3 - 2.0 * c1 - 4.0 * c2 * t) / ss;
        return (342.0 * v0 * u * u3 * u3 * u3 + 2.0 * c1 + 12.0 * c2 * t) / ss;
    }
    if (MODEL == PGT_POT_POLYLJ) {   // the reference's stiffness really uses 4 c2 t here (PolydisperseLJ.h:68)
        if (what == PGT_WANT_ENERGY) return v0 * u3 * u3 - v0 * u3 + c0 + c1 * t + c2 * t * t;
        if (what == PGT_WANT_FDIVR) return (12.0 * v0 * u * u3 * u3 - 6.0 * v0 * u * u3 - 2.0 * c1 - 4.0 * c2 * t) / ss;
        return (156.0 * v0 * u * u3 * u3 - 42.0 * v0 * u * u3 + 2.0 * c1 + 4.0 * c2 * t) / ss;
    }
    const double u5 = u * u * u * u * u;
    if (MODEL == PGT_POT_POLY10) {
        const double c3 = P.p[5];
        if (what == PGT_WANT_ENERGY) return v0 * u5 + c0 + c1 * t + c2 * t * t + c3 * t * t * t;
        if (what == PGT_WANT_FDIVR) return (10.0 * v0 * u * u5 - 2.0 * c1 - 4.0 * c2 * t - 6.0 * c3 * t * t) / ss;
        return (110.0 * v0 * u3 * u3 + 2.0 * c1 + 12.0 * c2 * t + 30.0 * c3 * t * t) / ss;
    }
    // PGT_POT_POLY106
    if (what == PGT_WANT_ENERGY) return v0 * (u5 - u * u * u) + c0 + c1 * t + c2 * t * t;
    if (what == PGT_WANT_FDIVR) return (v0 * (10.0 * u5 * u - 6.0 * u * u * u * u) - 2.0 * c1 - 4.0 * c2 * t) / ss;
    return (v0 * (110.0 * u3 * u3 - 42.0 * u * u * u * u) + 2.0 * c1 + 12.0 * c2 * t) / ss;
}

// Expands a runtime model id into the templated call.
#define PGT_DISPATCH_MODEL(model, CALL)                                      \
    switch (model) {                                                         \
    case PGT_POT_LJ: { constexpr int M = PGT_POT_LJ; CALL; } break;          \
    case PGT_POT_FORCESHIFTED_LJ: { constexpr int M = PGT_POT_FORCESHIFTED_LJ; CALL; } break; \
    case PGT_POT_POLY12: { constexpr int M = PGT_POT_POLY12; CALL; } break;  \
    case PGT_POT_POLY18: { constexpr int M = PGT_POT_POLY18; CALL; } break;  \
    case PGT_POT_POLY10: { constexpr int M = PGT_POT_POLY10; CALL; } break;  \
    case PGT_POT_POLY106: { constexpr int M = PGT_POT_POLY106; CALL; } break; \
    case PGT_POT_POLYYUKAWA: { constexpr int M = PGT_POT_POLYYUKAWA; CALL; } break; \
    case PGT_POT_POLYLJ: { constexpr int M = PGT_POT_POLYLJ; CALL; } break;  \
    default: break;                                                          \
    }

// --------------------------------------------------------------------------------------------
// Coarse-graining kernels Delta(d^2) = norm * shape(d^2/R^2) for d^2 < R^2, else 0.
struct PgtCG {
    int kernel;
    int order;
    double rcut;
    double rcutsq;
    double norm;        // octic 15/(8 pi R^2); rect 1/(pi R^2); mollifier 13.468.../(2 pi R^2)
};

PGT_HD double pgt_cg_norm(int kernel, double rcut)
{
    const double rcutsq = rcut * rcut;
    if (kernel == PGT_CG_OCTIC) return 15.0 / (8 * M_PI * rcutsq);
    if (kernel == PGT_CG_RECT) return 1 / (M_PI * rcutsq);
    return 13.46842098743079272180855772032086697297031125109170011102 / (2.0 * M_PI * rcut * rcut);
}

// Exact restatement of the reference's Delta for the scalar getters and the local (per-particle)
// observables; the bond-line quadrature uses pgt_cg_shape on the pre-scaled argument instead.
PGT_HD double pgt_cg_delta_exact(const PgtCG& cg, double dr_sq)
{
    if (!(dr_sq < cg.rcutsq)) return 0.0;
    if (cg.kernel == PGT_CG_OCTIC) {
        const double r4 = dr_sq * dr_sq / (cg.rcutsq * cg.rcutsq);
        const double r8 = r4 * r4;
        return cg.norm * (1 - 2 * r4 + r8);
    }
    if (cg.kernel == PGT_CG_RECT) return cg.norm;
    const double r2 = dr_sq / cg.rcutsq;
    return cg.norm * exp(-1.0 / (1 - r2));
}

// shape(t), t = d^2/R^2 in [0,1): octic (1 - t^2)^2, rect 1, mollifier exp(-1/(1-t)).
template <int KERNEL>
PGT_HD double pgt_cg_shape(double t)
{
    if (KERNEL == PGT_CG_OCTIC) {
        const double g = fma(-t, t, 1.0);
        return g * g;
    }
    if (KERNEL == PGT_CG_RECT) return 1.0;
    return exp(-1.0 / (1.0 - t));
}

// --------------------------------------------------------------------------------------------
// Adaptive bond quadrature: the reference's AdaptiveCGFunc (CoarseGrainFunction.h:94-134) calls GSLQuadrature
// (Quadrature.h:39-51) = gsl_integration_qag(f, 0, 1, epsabs, epsrel, limit, key 1, ...): global adaptive bisection
// with the 15-point Gauss-Kronrod pair, worst interval first.  The interval store here has a fixed capacity per thread
// (PGT_QAG_CAP); a bond that needs more (or any of GSL's failure exits, which abort the reference through GSL's
// default error handler) raises a status instead of a value nobody checked.
#define PGT_QAG_CAP 40
enum { PGT_QAG_OK = 0, PGT_QAG_ROUNDOFF0 = 1, PGT_QAG_LIMIT = 2, PGT_QAG_ROUNDOFF = 3, PGT_QAG_SINGULAR = 4, PGT_QAG_FAILED = 5,
       PGT_QAG_CAPACITY = 6 };

struct PgtGK15 { double r, err, rabs, rasc; };

// one Gauss-Kronrod (7, 15) evaluation on [a, b], sums in GSL's order (qk.c): Gauss nodes first, then the rest
#if defined(__CUDACC__)
#pragma nv_exec_check_disable
#endif
template <class F>
PGT_HD PgtGK15 pgt_gk15(F f, double a, double b)
{
    const double xgk[8] = { 0.991455371120812639206854697526329, 0.949107912342758524526189684047851,
                            0.864864423359769072789712788640926, 0.741531185599394439863864773280788,
                            0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
                            0.207784955007898467600689403773245, 0.0 };
    const double wg[4] = { 0.129484966168869693270611432679082, 0.279705391489276667901467771423780,
                           0.381830050505118944950369775488975, 0.417959183673469387755102040816327 };
    const double wgk[8] = { 0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
                            0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
                            0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
                            0.204432940075298892414161999234649, 0.209482141084727828012999174891714 };
    const double c = 0.5 * (a + b), h = 0.5 * (b - a);
    double lo[7], hi[7];
    const double fc = f(c);
    double gauss = fc * wg[3], kron = fc * wgk[7], sabs = fabs(kron);
PGT_UNROLL
    for (int j = 0; j < 3; ++j) {
        const int m = 2 * j + 1;
        const double d = h * xgk[m];
        lo[m] = f(c - d); hi[m] = f(c + d);
        const double fs = lo[m] + hi[m];
        gauss += wg[j] * fs;
        kron += wgk[m] * fs;
        sabs += wgk[m] * (fabs(lo[m]) + fabs(hi[m]));
    }
PGT_UNROLL
    for (int j = 0; j < 4; ++j) {
        const int m = 2 * j;
        const double d = h * xgk[m];
        lo[m] = f(c - d); hi[m] = f(c + d);
        kron += wgk[m] * (lo[m] + hi[m]);
        sabs += wgk[m] * (fabs(lo[m]) + fabs(hi[m]));
    }
    const double mean = kron * 0.5;
    double sasc = wgk[7] * fabs(fc - mean);
PGT_UNROLL
    for (int m = 0; m < 7; ++m) sasc += wgk[m] * (fabs(lo[m] - mean) + fabs(hi[m] - mean));
    PgtGK15 o;
    double err = fabs((kron - gauss) * h);
    o.r = kron * h; o.rabs = sabs * fabs(h); o.rasc = sasc * fabs(h);
    if (o.rasc != 0 && err != 0) {                       // rescale_error (err.c)
        const double scale = pow(200 * err / o.rasc, 1.5);
        err = scale < 1 ? o.rasc * scale : o.rasc;
    }
    if (o.rabs > 2.2250738585072014e-308 / (50 * 2.2204460492503131e-16)) {
        const double floor_err = 50 * 2.2204460492503131e-16 * o.rabs;
        if (floor_err > err) err = floor_err;
    }
    o.err = err;
    return o;
}

#if defined(__CUDACC__)
#pragma nv_exec_check_disable
#endif
template <class F>
PGT_HD double pgt_qag15(F f, double a, double b, double epsabs, double epsrel, int limit, int* status)
{
    const double EPS = 2.2204460492503131e-16, TINY = 2.2250738585072014e-308;
    double ia[PGT_QAG_CAP], ib[PGT_QAG_CAP], ir[PGT_QAG_CAP], ie[PGT_QAG_CAP];
    const PgtGK15 first = pgt_gk15(f, a, b);
    double tol = fmax(epsabs, epsrel * fabs(first.r));
    *status = PGT_QAG_OK;
    if (first.err <= 50 * EPS * first.rabs && first.err > tol) { *status = PGT_QAG_ROUNDOFF0; return first.r; }
    if ((first.err <= tol && first.err != first.rasc) || first.err == 0.0) return first.r;
    if (limit == 1) { *status = PGT_QAG_LIMIT; return first.r; }
    ia[0] = a; ib[0] = b; ir[0] = first.r; ie[0] = first.err;
    int n = 1, iter = 1, ro1 = 0, ro2 = 0, bad = 0;
    double area = first.r, errsum = first.err;
    do {
        if (n == PGT_QAG_CAP) { *status = PGT_QAG_CAPACITY; break; }
        int w = 0;                                       // worst interval, first of equals
        for (int k = 1; k < n; ++k) if (ie[k] > ie[w]) w = k;
        const double wa = ia[w], wb = ib[w], wr = ir[w], we = ie[w], mid = 0.5 * (wa + wb);
        const PgtGK15 L = pgt_gk15(f, wa, mid), R = pgt_gk15(f, mid, wb);
        const double area12 = L.r + R.r, err12 = L.err + R.err;
        errsum += (err12 - we);
        area += area12 - wr;
        if (L.rasc != L.err && R.rasc != R.err) {
            if (fabs(wr - area12) <= 1.0e-5 * fabs(area12) && err12 >= 0.99 * we) ++ro1;
            if (iter >= 10 && err12 > we) ++ro2;
        }
        tol = fmax(epsabs, epsrel * fabs(area));
        if (errsum > tol) {
            if (ro1 >= 6 || ro2 >= 20) bad = 2;
            const double small = (1 + 100 * EPS) * (fabs(mid) + 1000 * TINY);
            if (fabs(wa) <= small && fabs(wb) <= small) bad = 3;
        }
        if (R.err > L.err) {                             // the worse half keeps the slot, the other is appended
            ia[w] = mid; ir[w] = R.r; ie[w] = R.err;
            ia[n] = wa; ib[n] = mid; ir[n] = L.r; ie[n] = L.err;
        } else {
            ib[w] = mid; ir[w] = L.r; ie[w] = L.err;
            ia[n] = mid; ib[n] = wb; ir[n] = R.r; ie[n] = R.err;
        }
        ++n;
        ++iter;
    } while (iter < limit && !bad && errsum > tol);
    double sum = 0.0;
    for (int k = 0; k < n; ++k) sum += ir[k];
    if (*status == PGT_QAG_CAPACITY) return sum;
    if (errsum <= tol) *status = PGT_QAG_OK;
    else if (bad == 2) *status = PGT_QAG_ROUNDOFF;
    else if (bad == 3) *status = PGT_QAG_SINGULAR;
    else if (iter == limit) *status = PGT_QAG_LIMIT;
    else *status = PGT_QAG_FAILED;
    return sum;
}

