// ref_loops_shim.cpp -- C-ABI shim over the REFERENCE'S OWN calculators (test infrastructure only).
//
// Compiled by oracle/Makefile (target `ref`) with -I/root/reference -I/root/reference/pyglasstools, so that
//   irvingkirkwood/IrvingKirkwood.h (IrvingKirkwood::compute, :74-110), thermo/ThermoCalculator.h (:68-90),
//   ListofCommonObservables.h (:16-197), irvingkirkwood/cgfunc/CoarseGrainFunction.h (:56-134),
//   cgfunc/FixedPointQuadrature.h, cgfunc/Quadrature.h, irvingkirkwood/CoarseGrainedField.h, thermo/ThermoProperty.h,
//   Calculator.h, ParticleSystem.{h,cc}, SimBox.{h,cc}, potential/*.h
// are the reference's code, compiled where it lies; nothing is copied into this repo.  Three stand-ins under
// oracle/stubs/ replace the dependencies this image lacks: Aboria.h (particle container + euclidean_search, brute
// force), gsl/gsl_integration.h (glfixed per GSL's published rule, qag handed to the oracle's restatement) and the MPI
// log-file classes (MPIFile.h is skipped through its include guard; no output is written here).
// What this pins: the loop nests, the functors, the observables' accumulation and normalisation, max_rcut, the box.
// What stays restated (unpinned): Aboria's search and domain enforcement, GSL's rules.
// Output: oracle/_ref/libpgt_refloops.so (git-ignored, travels with gpurun).
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#include <omp.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <pybind11/eigen.h>
#include <Eigen/Dense>
#include <Eigen/StdVector>
#include <Eigen/Core>
#include <Eigen/CXX11/Tensor>
#include <Aboria.h>
#include <gsl/gsl_errno.h>
#include <gsl/gsl_integration.h>

// ---- MPI stand-in: MPIFile.h / MPIInterface.h need <mpi.h>, cereal and the vendored pybind11 2.4 (which does not
// compile against Python 3.12); the calculators only name these two classes
#define __MPI_FILE_H__
namespace MPI {
class Communicator {};
class LogFile {
public:
    void write_shared(std::string) {}
};
}

// the observables keep their values private and the reference has no numeric getter (only 6-digit text)
#define private public
#define protected public
#include <pyglasstools/SimBox.cc>
#include <pyglasstools/ParticleSystem.cc>
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
#include <pyglasstools/irvingkirkwood/cgfunc/CoarseGrainFunction.h>
#include <pyglasstools/ListofCommonObservables.h>
#include <pyglasstools/irvingkirkwood/CoarseGrainedField.h>
#include <pyglasstools/irvingkirkwood/IrvingKirkwood.h>
#include <pyglasstools/thermo/ThermoProperty.h>
#include <pyglasstools/thermo/ThermoCalculator.h>
#undef private
#undef protected

// typedefs of module-irvingkirkwood.cpp:8-13 and module-thermo.cpp:7-20
typedef ForceField<2, 2, VirialStress> VirialStressField2D;
typedef ForceField<2, 2, ElasticVirialStress> ElasticVirialStressField2D;
typedef LocalField<2, 2, KineticStress> KineticStressField2D;
typedef ForceScalarProperty<2, PotentialEnergy> GlobalPotentialEnergy2D;
typedef ForceProperty<2, 2, VirialStress> GlobalVirialStress2D;
typedef ForceProperty<2, 2, ElasticVirialStress> GlobalElasticVirialStress2D;
typedef LocalProperty<2, 2, KineticStress> GlobalKineticStress2D;
typedef ForceProperty<4, 2, BornStiffnessTensor> GlobalBornTensor2D;
typedef ForceField<2, 3, VirialStress> VirialStressField3D;
typedef ForceField<2, 3, ElasticVirialStress> ElasticVirialStressField3D;
typedef LocalField<2, 3, KineticStress> KineticStressField3D;
typedef ForceScalarProperty<3, PotentialEnergy> GlobalPotentialEnergy3D;
typedef ForceProperty<2, 3, VirialStress> GlobalVirialStress3D;
typedef ForceProperty<2, 3, ElasticVirialStress> GlobalElasticVirialStress3D;
typedef LocalProperty<2, 3, KineticStress> GlobalKineticStress3D;
typedef ForceProperty<4, 3, BornStiffnessTensor> GlobalBornTensor3D;

// ---- GSL stand-in, implemented on the oracle's tables / restatement (liborc.so)
extern "C" int orc_gl_table(int n, double* x, double* w);
extern "C" int orc_gsl_qag15(double (*f)(double, void*), void* params, double a, double b, double epsabs, double epsrel, int limit,
                             double* result, double* abserr);

extern "C" gsl_integration_glfixed_table* gsl_integration_glfixed_table_alloc(size_t n)
{
    gsl_integration_glfixed_table* t = new gsl_integration_glfixed_table;
    const size_t m = (n + 1) >> 1;
    t->n = n;
    t->x = new double[m];
    t->w = new double[m];
    t->precomputed = 0;
    orc_gl_table((int)n, t->x, t->w);
    return t;
}
extern "C" void gsl_integration_glfixed_table_free(gsl_integration_glfixed_table* t)
{
    if (!t) return;
    delete[] t->x;
    delete[] t->w;
    delete t;
}
extern "C" double gsl_integration_glfixed(const gsl_function* f, double a, double b, const gsl_integration_glfixed_table* t)
{
    const double* const x = t->x;
    const double* const w = t->w;
    const int n = (int)t->n, m = (n + 1) >> 1;
    const double A = 0.5 * (b - a), B = 0.5 * (b + a);
    double s;
    if (n & 1) {
        s = w[0] * f->function(B, f->params);
        for (int i = 1; i < m; i++) {
            const double Ax = A * x[i];
            s += w[i] * (f->function(B + Ax, f->params) + f->function(B - Ax, f->params));
        }
    } else {
        s = 0.0;
        for (int i = 0; i < m; i++) {
            const double Ax = A * x[i];
            s += w[i] * (f->function(B + Ax, f->params) + f->function(B - Ax, f->params));
        }
    }
    return A * s;
}
extern "C" gsl_integration_workspace* gsl_integration_workspace_alloc(size_t n)
{
    gsl_integration_workspace* w = new gsl_integration_workspace;
    w->limit = n;
    return w;
}
extern "C" void gsl_integration_workspace_free(gsl_integration_workspace* w) { delete w; }
extern "C" int gsl_integration_qag(const gsl_function* f, double a, double b, double epsabs, double epsrel, size_t limit, int key,
                                   gsl_integration_workspace*, double* result, double* abserr)
{
    if (key != 1) return -1;                               // Quadrature.h:47-49 always asks for the 15-point rule
    return orc_gsl_qag15(f->function, f->params, a, b, epsabs, epsrel, (int)limit, result, abserr);
}

// ---- builders
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

// cgfunc/__init__.py:27-82 -> module-cgfunc.cpp:23-30
static std::shared_ptr<CoarseGrainFunction> make_cg(int kernel, double rcut, int order, int adaptive, double epsrel, double epsabs, int maxiter)
{
    if (!adaptive) {
        switch (kernel) {
        case 0: return std::make_shared<FixedPointCGFunc<Octic>>(order, rcut);
        case 1: return std::make_shared<FixedPointCGFunc<Rect>>(order, rcut);
        case 2: return std::make_shared<FixedPointCGFunc<Mollifier>>(order, rcut);
        }
    } else {
        switch (kernel) {
        case 0: return std::make_shared<AdaptiveCGFunc<Octic>>(maxiter, epsrel, epsabs, rcut);
        case 1: return std::make_shared<AdaptiveCGFunc<Rect>>(maxiter, epsrel, epsabs, rcut);
        case 2: return std::make_shared<AdaptiveCGFunc<Mollifier>>(maxiter, epsrel, epsabs, rcut);
        }
    }
    return nullptr;
}

// utils.py:9-22 (simbox: Lz = 1 in 2-D, origin 0) and utils.py:41-69 (ParticleSystem from the frame's arrays)
static std::shared_ptr<ParticleSystem> make_system(const double* box, int64_t n, const double* pos, const double* vel,
                                                   const double* diam, const double* mass, const double* disp, int dim = 2)
{
    Eigen::Vector3d L(box[0], box[1], dim == 2 ? 1.0 : box[2]), origin(0, 0, 0);
    auto sb = std::make_shared<SimBox>(L, origin, dim);
    std::vector<double> d(diam, diam + n), m(n, 1.0);
    if (mass) m.assign(mass, mass + n);
    std::vector<std::vector<double>> r((size_t)n, std::vector<double>(3)), v((size_t)n, std::vector<double>(3, 0.0));
    for (int64_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) r[i][k] = pos[3 * i + k];
        if (vel) for (int k = 0; k < 3; ++k) v[i][k] = vel[3 * i + k];
    }
    auto sys = std::make_shared<ParticleSystem>(sb, (unsigned)n, d, m, r, v);
    if (disp) {
        if ((int64_t)sys->particles.size() != n) return nullptr;     // setDisplacement indexes the surviving particles
        std::vector<std::vector<double>> u((size_t)n, std::vector<double>(3));
        for (int64_t i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) u[i][k] = disp[3 * i + k];
        sys->setDisplacement(u);
    }
    return sys;
}

extern "C" {

// IrvingKirkwood::compute on G grid points with the reference's own field observables.
// obs_mask: bit0 VirialStressField2D, bit1 KineticStressField2D (the reference's order-dependent compute_cg),
// bit3 ElasticVirialStressField2D.  Outputs (NULL to skip): G x 4 doubles, col-major 2x2 per point.
int ref_ik_compute(const double* box, int64_t n, const double* pos, const double* vel, const double* diam, const double* mass,
                   const double* disp, int model, double scaled_rcut, const double* params, int nparams, int kernel, double cg_rcut,
                   int order, int adaptive, double epsrel, double epsabs, int maxiter, int64_t G, const double* grid,
                   unsigned obs_mask, double* virial, double* kinetic, double* elastic, double* max_rcut_out, int64_t* n_live)
{
    auto sys = make_system(box, n, pos, vel, diam, mass, disp);
    auto pot = make_pot(model, scaled_rcut, params, nparams);
    auto cg = make_cg(kernel, cg_rcut, order, adaptive, epsrel, epsabs, maxiter);
    if (!sys || !pot || !cg) return -1;
    if (n_live) *n_live = (int64_t)sys->particles.size();
    IrvingKirkwood calc(sys, pot, cg, std::make_shared<MPI::Communicator>());
    if (max_rcut_out) *max_rcut_out = calc.max_rcut;
    std::vector<Eigen::Vector3d> pts((size_t)G);
    for (int64_t g = 0; g < G; ++g) pts[g] = Eigen::Vector3d(grid[3 * g], grid[3 * g + 1], grid[3 * g + 2]);
    // irvingkirkwood/__init__.py:8-24 (initialize_field)
    std::shared_ptr<VirialStressField2D> fv;
    std::shared_ptr<KineticStressField2D> fk;
    std::shared_ptr<ElasticVirialStressField2D> fe;
    if (obs_mask & 1u) { fv = std::make_shared<VirialStressField2D>("virialstress", "2-TENSOR", false, (int)G); calc.addObservable(fv); }
    if (obs_mask & 2u) { fk = std::make_shared<KineticStressField2D>("kineticstress", "2-TENSOR", true, (int)G); calc.addObservable(fk); }
    if (obs_mask & 8u) { fe = std::make_shared<ElasticVirialStressField2D>("elasticvirialstress", "2-TENSOR", false, (int)G); calc.addObservable(fe); }
    calc.compute(pts);
    for (int64_t g = 0; g < G; ++g)
        for (int q = 0; q < 4; ++q) {
            if (fv && virial) virial[4 * g + q] = fv->val[g].data()[q];
            // LocalField never sizes its tensors (CoarseGrainedField.h:58-64: resize of the vector only), so the reference's
            // kinetic field is a list of EMPTY tensors; nothing to read then
            if (fk && kinetic && fk->val[g].size() == 4) kinetic[4 * g + q] = fk->val[g].data()[q];
            if (fe && elastic) elastic[4 * g + q] = fe->val[g].data()[q];
        }
    return 0;
}

// ThermoCalculator::compute with the reference's own global observables (thermo/__init__.py:8-36).
// out: virial[4] | kinetic[4] | energy | born[16] | elastic[4]   (flat data() order of each Eigen::Tensor)
int ref_thermo_compute(const double* box, int64_t n, const double* pos, const double* vel, const double* diam, const double* mass,
                       const double* disp, int model, double scaled_rcut, const double* params, int nparams, unsigned obs_mask,
                       double* out29, double* max_rcut_out)
{
    auto sys = make_system(box, n, pos, vel, diam, mass, disp);
    auto pot = make_pot(model, scaled_rcut, params, nparams);
    if (!sys || !pot) return -1;
    ThermoCalculator calc(sys, pot);
    if (max_rcut_out) *max_rcut_out = calc.max_rcut;
    auto v = std::make_shared<GlobalVirialStress2D>("virialstress", "2-TENSOR", false);
    auto k = std::make_shared<GlobalKineticStress2D>("kineticstress", "2-TENSOR", true);
    auto e = std::make_shared<GlobalPotentialEnergy2D>("potentialenergy", "SCALAR", false);
    auto b = std::make_shared<GlobalBornTensor2D>("borntensor", "4-TENSOR", false);
    auto el = std::make_shared<GlobalElasticVirialStress2D>("elasticvirialstress", "2-TENSOR", false);
    if (obs_mask & 1u) calc.addObservable(v);
    if (obs_mask & 2u) calc.addObservable(k);
    if (obs_mask & 4u) calc.addObservable(e);
    if (obs_mask & 8u) calc.addObservable(b);
    if (obs_mask & 16u) calc.addObservable(el);
    calc.compute();
    for (int q = 0; q < 29; ++q) out29[q] = 0.0;
    if (obs_mask & 1u) for (int q = 0; q < 4; ++q) out29[q] = v->val.data()[q];
    if (obs_mask & 2u) for (int q = 0; q < 4; ++q) out29[4 + q] = k->val.data()[q];
    if (obs_mask & 4u) out29[8] = e->val;
    if (obs_mask & 8u) for (int q = 0; q < 16; ++q) out29[9 + q] = b->val.data()[q];
    if (obs_mask & 16u) for (int q = 0; q < 4; ++q) out29[25 + q] = el->val.data()[q];
    return 0;
}

// The same two calculators with the reference's Dim = 3 typedefs (module-irvingkirkwood.cpp:9-13, module-thermo.cpp:8-20); the box is
// periodic on all three axes (SimBox.cc:33, :41-44).  Field outputs: G x 9 (flat a + 3 b).
int ref_ik_compute3(const double* box, int64_t n, const double* pos, const double* vel, const double* diam, const double* mass,
                    const double* disp, int model, double scaled_rcut, const double* params, int nparams, int kernel, double cg_rcut,
                    int order, int adaptive, double epsrel, double epsabs, int maxiter, int64_t G, const double* grid,
                    unsigned obs_mask, double* virial, double* kinetic, double* elastic, double* max_rcut_out, int64_t* n_live)
{
    auto sys = make_system(box, n, pos, vel, diam, mass, disp, 3);
    auto pot = make_pot(model, scaled_rcut, params, nparams);
    auto cg = make_cg(kernel, cg_rcut, order, adaptive, epsrel, epsabs, maxiter);
    if (!sys || !pot || !cg) return -1;
    if (n_live) *n_live = (int64_t)sys->particles.size();
    IrvingKirkwood calc(sys, pot, cg, std::make_shared<MPI::Communicator>());
    if (max_rcut_out) *max_rcut_out = calc.max_rcut;
    std::vector<Eigen::Vector3d> pts((size_t)G);
    for (int64_t g = 0; g < G; ++g) pts[g] = Eigen::Vector3d(grid[3 * g], grid[3 * g + 1], grid[3 * g + 2]);
    std::shared_ptr<VirialStressField3D> fv;
    std::shared_ptr<KineticStressField3D> fk;
    std::shared_ptr<ElasticVirialStressField3D> fe;
    if (obs_mask & 1u) { fv = std::make_shared<VirialStressField3D>("virialstress", "2-TENSOR", false, (int)G); calc.addObservable(fv); }
    if (obs_mask & 2u) { fk = std::make_shared<KineticStressField3D>("kineticstress", "2-TENSOR", true, (int)G); calc.addObservable(fk); }
    if (obs_mask & 8u) { fe = std::make_shared<ElasticVirialStressField3D>("elasticvirialstress", "2-TENSOR", false, (int)G); calc.addObservable(fe); }
    calc.compute(pts);
    for (int64_t g = 0; g < G; ++g)
        for (int q = 0; q < 9; ++q) {
            if (fv && virial) virial[9 * g + q] = fv->val[g].data()[q];
            if (fk && kinetic && fk->val[g].size() == 9) kinetic[9 * g + q] = fk->val[g].data()[q];
            if (fe && elastic) elastic[9 * g + q] = fe->val[g].data()[q];
        }
    return 0;
}

// out: virial[9] | kinetic[9] | energy | born[81] | elastic[9]   (109 doubles)
int ref_thermo_compute3(const double* box, int64_t n, const double* pos, const double* vel, const double* diam, const double* mass,
                        const double* disp, int model, double scaled_rcut, const double* params, int nparams, unsigned obs_mask,
                        double* out109, double* max_rcut_out)
{
    auto sys = make_system(box, n, pos, vel, diam, mass, disp, 3);
    auto pot = make_pot(model, scaled_rcut, params, nparams);
    if (!sys || !pot) return -1;
    ThermoCalculator calc(sys, pot);
    if (max_rcut_out) *max_rcut_out = calc.max_rcut;
    auto v = std::make_shared<GlobalVirialStress3D>("virialstress", "2-TENSOR", false);
    auto k = std::make_shared<GlobalKineticStress3D>("kineticstress", "2-TENSOR", true);
    auto e = std::make_shared<GlobalPotentialEnergy3D>("potentialenergy", "SCALAR", false);
    auto b = std::make_shared<GlobalBornTensor3D>("borntensor", "4-TENSOR", false);
    auto el = std::make_shared<GlobalElasticVirialStress3D>("elasticvirialstress", "2-TENSOR", false);
    if (obs_mask & 1u) calc.addObservable(v);
    if (obs_mask & 2u) calc.addObservable(k);
    if (obs_mask & 4u) calc.addObservable(e);
    if (obs_mask & 8u) calc.addObservable(b);
    if (obs_mask & 16u) calc.addObservable(el);
    calc.compute();
    for (int q = 0; q < 109; ++q) out109[q] = 0.0;
    if (obs_mask & 1u) for (int q = 0; q < 9; ++q) out109[q] = v->val.data()[q];
    if (obs_mask & 2u) for (int q = 0; q < 9; ++q) out109[9 + q] = k->val.data()[q];
    if (obs_mask & 4u) out109[18] = e->val;
    if (obs_mask & 8u) for (int q = 0; q < 81; ++q) out109[19 + q] = b->val.data()[q];
    if (obs_mask & 16u) for (int q = 0; q < 9; ++q) out109[100 + q] = el->val.data()[q];
    return 0;
}

// CoarseGrainFunction::getDeltaFunc / getBondFunc through the reference's classes (for the quadrature-order tests)
double ref_cg_bond(int kernel, double cg_rcut, int order, int adaptive, double epsrel, double epsabs, int maxiter, const double* x,
                   const double* ri, const double* rij)
{
    auto cg = make_cg(kernel, cg_rcut, order, adaptive, epsrel, epsabs, maxiter);
    return cg->getBondFunc(Eigen::Vector3d(x[0], x[1], x[2]), Eigen::Vector3d(ri[0], ri[1], ri[2]), Eigen::Vector3d(rij[0], rij[1], rij[2]));
}

// SimBox as utils.simbox builds it: lower | upper | boxsize | periodic | vol
void ref_simbox(const double* box, int dim, double* out13)
{
    Eigen::Vector3d L(box[0], box[1], dim == 2 ? 1.0 : box[2]), origin(0, 0, 0);
    SimBox sb(L, origin, dim);
    for (int d = 0; d < 3; ++d) { out13[d] = sb.lowerbound[d]; out13[3 + d] = sb.upperbound[d]; out13[6 + d] = sb.boxsize[d]; out13[9 + d] = sb.periodic[d]; }
    out13[12] = sb.vol;
}

}
