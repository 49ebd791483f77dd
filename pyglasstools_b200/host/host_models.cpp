// host_models.cpp -- pair potentials and coarse-graining functions as Python-visible parameter holders.
// The scalar getters call the host-side evaluators exported by libpgt_cuda.so (pgt_pair_eval, pgt_cg_delta,
// pgt_cg_bond), i.e. the same source (pgt_math.cuh) the device kernels are compiled from.
#include <cmath>
#include <cstdio>
#include <stdexcept>
#include <string>

#include "host_common.h"

namespace pgt_host {

double PairPotential::eval(const Vec3& rij, double di, double dj, int what) const
{
    if (model < 0) return 0.0;       // abstract base: PairPotential.h:36-63 returns 0 everywhere
    return pgt_pair_eval(model, scaled_rcut, params.data(), (int)params.size(), rij.data(), di, dj, what);
}

std::string PairPotential::signature() const
{
    std::string s = std::to_string(model);
    char buf[40];
    std::snprintf(buf, sizeof buf, "|%a", scaled_rcut);
    s += buf;
    for (double p : params) {
        std::snprintf(buf, sizeof buf, "|%a", p);
        s += buf;
    }
    return s;
}

CoarseGrainFunction::CoarseGrainFunction(int kernel, int quad_mode, int order, double rcut)
    : kernel(kernel), quad_mode(quad_mode), order(order), cg_rcut(rcut), rcut_at_ctor(rcut)
{
}

// Delta(|x - ri|^2); CGFuncMollifier keeps the normalisation of the construction radius after setRcut
// (CGFuncMollifier.h:12,17), which the ratio below reproduces.
static double frozen_norm_ratio(const CoarseGrainFunction& f)
{
    if (f.kernel != PGT_CG_MOLLIFIER || f.cg_rcut == f.rcut_at_ctor) return 1.0;
    return (f.cg_rcut * f.cg_rcut) / (f.rcut_at_ctor * f.rcut_at_ctor);
}

double CoarseGrainFunction::getDeltaFunc(const Vec3& x, const Vec3& ri) const
{
    if (kernel < 0) return 0.0;
    double d2 = 0.0;
    for (int k = 0; k < 3; ++k) d2 += (x[k] - ri[k]) * (x[k] - ri[k]);
    return pgt_cg_delta(kernel, cg_rcut, d2) * frozen_norm_ratio(*this);
}

double CoarseGrainFunction::getObjFunc(double s, const Vec3& x, const Vec3& ri, const Vec3& rij) const
{
    const Vec3 r{ ri[0] + s * rij[0], ri[1] + s * rij[1], ri[2] + s * rij[2] };
    return getDeltaFunc(x, r);
}

double CoarseGrainFunction::getBondFunc(const Vec3& x, const Vec3& ri, const Vec3& rij) const
{
    if (kernel < 0) return 0.0;
    if (quad_mode != PGT_QUAD_FIXEDPOINT) {
        int status = 0;
        const double v = pgt_cg_bond_adaptive(kernel, cg_rcut, abserr, relerr, maxiter, x.data(), ri.data(), rij.data(), &status);
        // GSL's default error handler aborts the reference on any of these exits (Quadrature.h:46-48)
        if (status) throw std::runtime_error("[ERROR] gsl_integration_qag: tolerance not reached (exit code " + std::to_string(status) + ")");
        return v * frozen_norm_ratio(*this);
    }
    return pgt_cg_bond(kernel, cg_rcut, order, x.data(), ri.data(), rij.data()) * frozen_norm_ratio(*this);
}

namespace {

template <int MODEL>
struct PotT : PairPotential {
    explicit PotT(std::vector<double> p) : PairPotential(MODEL, 0.0, std::move(p)) {}
    PotT(double rc, std::vector<double> p) : PairPotential(MODEL, rc, std::move(p)) {}
};

template <int MODEL>
void export_pot(py::module_& m, const char* name)
{
    using T = PotT<MODEL>;
    py::class_<T, PairPotential, std::shared_ptr<T>>(m, name)
        .def(py::init<std::vector<double>>())
        .def(py::init<double, std::vector<double>>());
}

template <int KERNEL>
struct FixedCG : CoarseGrainFunction {
    FixedCG(int order, double rcut) : CoarseGrainFunction(KERNEL, PGT_QUAD_FIXEDPOINT, order, rcut)
    {
        if (order < 1) throw std::invalid_argument("[ERROR] quadrature order must be >= 1");
    }
};

template <int KERNEL>
struct AdaptiveCG : CoarseGrainFunction {
    AdaptiveCG(int maxiter_, double relerr_, double abserr_, double rcut) : CoarseGrainFunction(KERNEL, PGT_QUAD_ADAPTIVE, 0, rcut)
    {
        maxiter = maxiter_;
        relerr = relerr_;
        abserr = abserr_;
    }
};

template <class T, class... Args>
void export_cg(py::module_& m, const char* name)
{
    py::class_<T, CoarseGrainFunction, std::shared_ptr<T>>(m, name).def(py::init<Args...>());
}

}  // namespace

void export_models(py::module_& m)
{
    py::class_<PairPotential, std::shared_ptr<PairPotential>>(m, "PairPotential")
        .def(py::init([](std::vector<double> p) { return std::make_shared<PairPotential>(-1, 0.0, std::move(p)); }))
        .def(py::init([](double rc, std::vector<double> p) { return std::make_shared<PairPotential>(-1, rc, std::move(p)); }))
        .def("getRcut", &PairPotential::getRcut)
        .def("getPairForceDivR", &PairPotential::getPairForceDivR)
        .def("getPairEnergy", &PairPotential::getPairEnergy)
        .def("getBondStiffness", &PairPotential::getBondStiffness)
        .def("getParams", &PairPotential::getParams)
        .def("setParams", &PairPotential::setParams)
        .def_readwrite("scaled_rcut", &PairPotential::scaled_rcut);
    export_pot<PGT_POT_LJ>(m, "PairPotentialLJ");
    export_pot<PGT_POT_FORCESHIFTED_LJ>(m, "PairPotentialForceShiftedLJ");
    export_pot<PGT_POT_POLY12>(m, "PairPotentialPoly12");
    export_pot<PGT_POT_POLY18>(m, "PairPotentialPoly18");
    export_pot<PGT_POT_POLY10>(m, "PairPotentialPoly10");
    export_pot<PGT_POT_POLY106>(m, "PairPotentialPoly106");
    export_pot<PGT_POT_POLYYUKAWA>(m, "PairPotentialPolyYukawa");
    export_pot<PGT_POT_POLYLJ>(m, "PairPotentialPolyLJ");

    py::class_<CoarseGrainFunction, std::shared_ptr<CoarseGrainFunction>>(m, "CoarseGrainFunction")
        .def(py::init([](double rcut) { return std::make_shared<CoarseGrainFunction>(-1, PGT_QUAD_FIXEDPOINT, 0, rcut); }))
        .def("getRcut", &CoarseGrainFunction::getRcut)
        .def("setRcut", &CoarseGrainFunction::setRcut)
        .def("getDeltaFunc", &CoarseGrainFunction::getDeltaFunc)
        .def("getObjFunc", &CoarseGrainFunction::getObjFunc)
        .def("getBondFunc", &CoarseGrainFunction::getBondFunc);
    export_cg<FixedCG<PGT_CG_OCTIC>, int, double>(m, "CGFuncOcticFixedPoint");
    export_cg<FixedCG<PGT_CG_RECT>, int, double>(m, "CGFuncRectFixedPoint");
    export_cg<FixedCG<PGT_CG_MOLLIFIER>, int, double>(m, "CGFuncMollifierFixedPoint");
    export_cg<AdaptiveCG<PGT_CG_OCTIC>, int, double, double, double>(m, "CGFuncOcticAdaptive");
    export_cg<AdaptiveCG<PGT_CG_RECT>, int, double, double, double>(m, "CGFuncRectAdaptive");
    export_cg<AdaptiveCG<PGT_CG_MOLLIFIER>, int, double, double, double>(m, "CGFuncMollifierAdaptive");
}

}  // namespace pgt_host
