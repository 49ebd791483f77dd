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
