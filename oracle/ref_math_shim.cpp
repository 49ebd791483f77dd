// ref_math_shim.cpp -- C-ABI shim over the REFERENCE'S OWN headers (test infrastructure only).
//
// Compiled by oracle/Makefile with -I/root/reference so that the pair models
// (pyglasstools/potential/PairPotential.h + the model headers listed in
// module-potential.cpp:1-8) and the coarse-graining kernels
// (pyglasstools/irvingkirkwood/cgfunc/CGFunc{Octic,Rect,Mollifier}.h) are the
// reference's code, compiled where it lies; nothing is copied into this repo.
// Output goes to oracle/_ref/libpgt_refmath.so (git-ignored, travels with gpurun).
// The rest of the reference path (Aboria search, GSL quadrature, MPI) cannot be
// built here: Boost, GSL and MPI are absent ("Oracle").
#include <pyglasstools/potential/PairPotential.h>
#include <pyglasstools/potential/LennardJones.h>
#include <pyglasstools/potential/Polydisperse12.h>
#include <pyglasstools/potential/Polydisperse18.h>
#include <pyglasstools/potential/PolydisperseLJ.h>
#include <pyglasstools/potential/Polydisperse10.h>
#include <pyglasstools/potential/Polydisperse106.h>
#include <pyglasstools/potential/PolydisperseYukawa.h>
#include <pyglasstools/irvingkirkwood/cgfunc/CGFuncOctic.h>
#include <pyglasstools/irvingkirkwood/cgfunc/CGFuncRect.h>
#include <pyglasstools/irvingkirkwood/cgfunc/CGFuncMollifier.h>
#include <memory>
#include <cstdint>

static std::shared_ptr<PairPotential> make_pot(int model, double rc, const double* p, int np)
{
    std::vector<double> v(p, p + np);
    switch (model) {   // ids follow oracle/pgt_oracle.cpp and module-potential.cpp:13-20
    case 0: return std::make_shared<ShortRangePairPotential<LennardJones>>(rc, v);
    case 1: return std::make_shared<ShortRangePairPotential<ForceShiftedLennardJones>>(rc, v);
    case 2: return std::make_shared<ShortRangePairPotential<Polydisperse12>>(rc, v);
    case 3: return std::make_shared<ShortRangePairPotential<Polydisperse18>>(rc, v);
    case 4: return std::make_shared<ShortRangePairPotential<Polydisperse10>>(rc, v);
    case 5: return std::make_shared<ShortRangePairPotential<Polydisperse106>>(rc, v);
    case 6: return std::make_shared<ShortRangePairPotential<PolydisperseYukawa>>(rc, v);
    case 7: return std::make_shared<ShortRangePairPotential<PolydisperseLJ>>(rc, v);
    }
    return nullptr;
}

extern "C" {

int ref_pair_table(int model, double scaled_rcut, const double* params, int nparams, int64_t n,
                   const double* rij, const double* di, const double* dj,
                   double* energy, double* fdivr, double* stiff, double* rcutsq)
{
    auto pot = make_pot(model, scaled_rcut, params, nparams);
    if (!pot) return -1;
    for (int64_t k = 0; k < n; ++k) {
        Eigen::Vector3d r(rij[3 * k], rij[3 * k + 1], rij[3 * k + 2]);
        energy[k] = pot->getPairEnergy(r, di[k], dj[k]);
        fdivr[k] = pot->getPairForceDivR(r, di[k], dj[k]);
        stiff[k] = pot->getBondStiffness(r, di[k], dj[k]);
        rcutsq[k] = pot->getRcut(r, di[k], dj[k]);
    }
    return 0;
}

int ref_cg_table(int kernel, double rcut, int64_t n, const double* dr_sq, double* out)
{
    Octic o(rcut); Rect r(rcut); Mollifier m(rcut);
    for (int64_t k = 0; k < n; ++k) {
        switch (kernel) {
        case 0: out[k] = o.compute(dr_sq[k]); break;
        case 1: out[k] = r.compute(dr_sq[k]); break;
        case 2: out[k] = m.compute(dr_sq[k]); break;
        default: return -1;
        }
    }
    return 0;
}

}
