// host_calc.cpp -- ThermoCalculator / IrvingKirkwood and their observable holders.  compute() is a
// sequence of C-ABI calls (upload -> cell list -> half-bonds -> gather / thermo sums -> download); the
// GIL is released while the device works.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>
#include <sstream>

#include "host_common.h"
#include "host_io.h"

namespace pgt_host {

// ------------------------------------------------------------------------------------ observables
static int ipow(int b, int e) { int r = 1; while (e-- > 0) r *= b; return r; }

ThermoObservable::ThermoObservable(ThermoKind k, int d, std::string n, std::string t, bool local)
    : Observable(std::move(n), std::move(t), local, d), kind(k)
{
    const int rank = (k == ThermoKind::Energy) ? 0 : (k == ThermoKind::Born ? 4 : 2);
    val.assign((std::size_t)ipow(d, rank), 0.0);
}

unsigned ThermoObservable::mask() const
{
    switch (kind) {
    case ThermoKind::Energy: return PGT_OBS_ENERGY;
    case ThermoKind::Virial: return PGT_OBS_VIRIAL;
    case ThermoKind::Kinetic: return PGT_OBS_KINETIC;
    case ThermoKind::Born: return PGT_OBS_BORN;
    default: return 0;
    }
}

void ThermoObservable::save(std::shared_ptr<LogFile> file, int index) const
{
    if (index < 0 || (std::size_t)index >= val.size()) throw std::invalid_argument("[ERROR] observable component index out of range");
    file->write_shared(std::to_string(val[(std::size_t)index]));      // ThermoProperty.h:76,120,196
}

py::array_t<double> ThermoObservable::values() const
{
    py::array_t<double> a((py::ssize_t)val.size());
    std::memcpy(a.mutable_data(), val.data(), val.size() * sizeof(double));
    return a;
}

FieldObservable::FieldObservable(FieldKind k, int d, std::string n, std::string t, bool local, int gridsize)
    : Observable(std::move(n), std::move(t), local, d), kind(k), ncomp(d * d), npoints((std::size_t)std::max(gridsize, 0))
{
    val.assign(npoints * (std::size_t)ncomp, 0.0);
}

void FieldObservable::addGridpoints(DArray pts)
{
    if (pts.ndim() != 2 || pts.shape(1) < 3) throw std::invalid_argument("[ERROR] grid points must be an (G,3) array");
    points.resize(3 * (std::size_t)pts.shape(0));
    auto r = pts.unchecked<2>();
    for (py::ssize_t i = 0; i < pts.shape(0); ++i)
        for (int k = 0; k < 3; ++k) points[3 * i + k] = r(i, k);
}

void FieldObservable::clear_point(int grid_id)
{
    if (grid_id < 0 || (std::size_t)grid_id >= npoints) throw std::invalid_argument("[ERROR] grid index out of range");
    std::fill_n(val.begin() + (std::size_t)grid_id * ncomp, ncomp, 0.0);
}

std::string FieldObservable::gettostring(int index, int grid_id) const
{
    if (grid_id < 0 || (std::size_t)grid_id >= npoints || index < 0 || index >= ncomp)
        throw std::invalid_argument("[ERROR] field index out of range");
    return ostream_double(val[(std::size_t)grid_id * ncomp + index]) + " ";   // CoarseGrainedField.h:160-183
}

void FieldObservable::save(std::shared_ptr<LogFile> file, int index, int grid_id) const
{
    file->write_shared(gettostring(index, grid_id));
}

py::array_t<double> FieldObservable::values() const
{
    py::array_t<double> a({ (py::ssize_t)npoints, (py::ssize_t)ncomp });
    if (!val.empty()) std::memcpy(a.mutable_data(), val.data(), val.size() * sizeof(double));
    return a;
}

// ------------------------------------------------------------------------------------ Calculator
Calculator::Calculator(std::shared_ptr<ParticleSystem> sys, std::shared_ptr<PairPotential> pot)
    : m_sysdata(std::move(sys)), m_potential(std::move(pot))
{
    if (!m_sysdata || !m_potential) throw std::invalid_argument("[ERROR] Calculator needs a particle system and a pair potential");
    if (m_sysdata->size() == 0) throw std::invalid_argument("[ERROR] Calculator: empty particle system");
    // Calculator.h:37-46 -- fixed here, never refreshed by setSystemData
    const double maxd = *std::max_element(m_sysdata->m_diam.begin(), m_sysdata->m_diam.end());
    max_rcut = std::max(m_potential->scaled_rcut * maxd, maxd);
}

void Calculator::prepare_device(bool need_bonds)
{
    DeviceSession& S = DeviceSession::get();
    pgt_ctx* ctx = S.ctx();
    const ParticleSystem& ps = *m_sysdata;
    const SimBox& box = *ps.simbox;
    if (box.dim != 2 && box.dim != 3) throw std::domain_error("[ERROR] the box dimension must be 2 or 3");
    if (box.origin[0] != 0.0 || box.origin[1] != 0.0 || (box.dim == 3 && box.origin[2] != 0.0))
        throw std::domain_error("[ERROR] a shifted box origin is not implemented on device");
    if (m_potential->model < 0) throw std::domain_error("[ERROR] the abstract PairPotential carries no force model");

    const std::string sig = m_potential->signature();
    if (sig != S.pot_signature) {
        S.check(pgt_potential_set(ctx, m_potential->model, m_potential->scaled_rcut, m_potential->params.data(),
                                  (int)m_potential->params.size()), "pgt_potential_set");
        S.pot_signature = sig;
        S.bonds_token = 0;
    }
    if (S.frame_token != ps.token() || S.frame_max_rcut != max_rcut) {
        S.check(pgt_set_max_rcut(ctx, max_rcut), "pgt_set_max_rcut");
        const double L[3] = { box.boxsize[0], box.boxsize[1], box.boxsize[2] };
        // a ParticleSystem refilled frame after frame keeps its arrays where they are: page-locked in place, the upload
        // runs at link speed (every compute path below ends with a synchronising read-back, so the arrays are not touched
        // again before the copy has completed)
        ps.pin_pos.ensure(ps.m_pos.data(), ps.m_pos.size() * sizeof(double));
        ps.pin_diam.ensure(ps.m_diam.data(), ps.m_diam.size() * sizeof(double));
        S.check(pgt_frame_upload(ctx, ps.m_pos.data(), ps.m_diam.data(), ps.unit_mass ? nullptr : ps.m_mass.data(),
                                 ps.has_velocity ? ps.m_vel.data() : nullptr, 0, (int64_t)ps.size(), L, box.dim), "pgt_frame_upload");
        S.check(pgt_build_cells(ctx), "pgt_build_cells");
        S.frame_token = ps.token();
        S.frame_max_rcut = max_rcut;
        S.bonds_token = 0;
    }
    if (need_bonds && S.bonds_token != ps.token()) {
        S.check(pgt_build_bonds(ctx), "pgt_build_bonds");
        S.bonds_token = ps.token();
    }
}

// ------------------------------------------------------------------------------------ thermo
void ThermoCalculator::compute()
{
    unsigned mask = 0;
    bool elastic = false;
    const int D = m_sysdata->simbox->dim, D2 = D * D, D4 = D2 * D2;
    for (auto& kv : m_obs) {
        // the reference's templates would let a 2-D observable read a 3-D system (and the reverse); nothing in its Python layer does
        if (kv.second->dim != D) throw std::domain_error("[ERROR] observable '" + kv.second->name + "' and the simulation box differ in dimension");
        if (kv.second->kind == ThermoKind::ElasticVirial) elastic = true;
        else mask |= kv.second->mask();
    }
    for (auto& kv : m_obs) kv.second->clear();
    if (mask == 0 && !elastic) return;
    double out[128] = { 0 }, out_el[9] = { 0 };
    {
        py::gil_scoped_release nogil;
        prepare_device(false);
        DeviceSession& S = DeviceSession::get();
        if (mask) S.check(pgt_thermo_compute(S.ctx(), mask, out), "pgt_thermo_compute");
        if (elastic) {
            // displacements may change between two computes on the same frame (set_displacement): always resend
            S.check(pgt_frame_set_displacement(S.ctx(), m_sysdata->m_disp.data(), (int64_t)m_sysdata->size()), "pgt_frame_set_displacement");
            S.check(pgt_thermo_elastic(S.ctx(), out_el), "pgt_thermo_elastic");
        }
    }
    // device layout: virial D^2 | kinetic D^2 | energy | born D^4, already times 1/vol (ThermoCalculator.h:58-64)
    for (auto& kv : m_obs) {
        ThermoObservable& o = *kv.second;
        switch (o.kind) {
        case ThermoKind::Virial: std::copy(out, out + D2, o.val.begin()); break;
        case ThermoKind::Kinetic: std::copy(out + D2, out + 2 * D2, o.val.begin()); break;
        case ThermoKind::Energy: o.val[0] = out[2 * D2]; break;
        case ThermoKind::Born: std::copy(out + 2 * D2 + 1, out + 2 * D2 + 1 + D4, o.val.begin()); break;
        case ThermoKind::ElasticVirial: std::copy(out_el, out_el + D2, o.val.begin()); break;
        default: break;
        }
    }
}

// ------------------------------------------------------------------------------------ Irving-Kirkwood
IrvingKirkwood::IrvingKirkwood(std::shared_ptr<ParticleSystem> sys, std::shared_ptr<PairPotential> pot,
                               std::shared_ptr<CoarseGrainFunction> cg, std::shared_ptr<Communicator> comm)
    : Calculator(std::move(sys), std::move(pot)), m_cg(std::move(cg)), m_comm(std::move(comm))
{
    if (!m_cg) throw std::invalid_argument("[ERROR] IrvingKirkwood needs a coarse-graining function");
}

void IrvingKirkwood::run(long long npoints)
{
    unsigned mask = 0;
    for (auto& kv : m_obs) {
        FieldObservable& o = *kv.second;
        if (o.dim != m_sysdata->simbox->dim) throw std::domain_error("[ERROR] field '" + o.name + "' and the simulation box differ in dimension");
        if ((long long)o.npoints != npoints)
            throw std::invalid_argument("[ERROR] field '" + o.name + "' was sized for " + std::to_string(o.npoints) + " grid points, compute() got " +
                                        std::to_string(npoints));
        mask |= (o.kind == FieldKind::Virial) ? PGT_OBS_VIRIAL : (o.kind == FieldKind::Kinetic ? PGT_OBS_KINETIC : PGT_OBS_ELASTIC);
    }
    if (mask == 0 || npoints == 0) return;
    if (m_cg->kernel < 0) throw std::domain_error("[ERROR] the abstract CoarseGrainFunction has no kernel");
    const double ratio = (m_cg->kernel == PGT_CG_MOLLIFIER) ? (m_cg->cg_rcut * m_cg->cg_rcut) / (m_cg->rcut_at_ctor * m_cg->rcut_at_ctor) : 1.0;

    py::gil_scoped_release nogil;
    DeviceSession& S = DeviceSession::get();
    pgt_ctx* ctx = S.ctx();
    if (m_cg->quad_mode == PGT_QUAD_ADAPTIVE)
        S.check(pgt_cg_set_adaptive(ctx, m_cg->kernel, m_cg->cg_rcut, m_cg->abserr, m_cg->relerr, m_cg->maxiter), "pgt_cg_set_adaptive");
    else
        S.check(pgt_cg_set(ctx, m_cg->kernel, m_cg->cg_rcut, m_cg->quad_mode, m_cg->order), "pgt_cg_set");
    prepare_device((mask & PGT_OBS_VIRIAL) != 0);
    if (mask & PGT_OBS_ELASTIC)
        S.check(pgt_frame_set_displacement(ctx, m_sysdata->m_disp.data(), (int64_t)m_sysdata->size()), "pgt_frame_set_displacement");
    S.check(pgt_ik_compute(ctx, mask), "pgt_ik_compute");
    for (auto& kv : m_obs) {
        FieldObservable& o = *kv.second;
        const int field = o.kind == FieldKind::Virial ? PGT_FIELD_VIRIAL : (o.kind == FieldKind::Kinetic ? PGT_FIELD_KINETIC : PGT_FIELD_ELASTIC);
        o.pin_val.ensure(o.val.data(), o.val.size() * sizeof(double));
        S.check(pgt_ik_download(ctx, field, o.val.data()), "pgt_ik_download");
        if (ratio != 1.0)
            for (double& v : o.val) v *= ratio;
    }
}

void IrvingKirkwood::compute_points(DArray gridpoints)
{
    if (gridpoints.ndim() != 2 || gridpoints.shape(1) != 3) throw std::invalid_argument("[ERROR] grid points must be an (G,3) array");
    const long long G = gridpoints.shape(0);
    DeviceSession& S = DeviceSession::get();
    S.check(pgt_grid_set_points(S.ctx(), gridpoints.data(), G), "pgt_grid_set_points");
    run(G);
}

void IrvingKirkwood::compute_cartesian(DArray xs, DArray ys, long long g0, long long g1)
{
    if (xs.ndim() != 1 || ys.ndim() != 1) throw std::invalid_argument("[ERROR] grid axes must be 1-D arrays");
    DeviceSession& S = DeviceSession::get();
    S.check(pgt_grid_set_cartesian(S.ctx(), xs.data(), (int)xs.shape(0), ys.data(), (int)ys.shape(0)), "pgt_grid_set_cartesian");
    S.check(pgt_grid_set_shard(S.ctx(), g0, g1), "pgt_grid_set_shard");
    run(g1 - g0);
}

void IrvingKirkwood::printDisplacement() const
{
    // IrvingKirkwood.h:66-72: one line per particle on rank 0
    if (m_comm && m_comm->rank() != 0) return;
    for (std::size_t i = 0; i < m_sysdata->size(); ++i)
        std::cout << m_sysdata->m_disp[3 * i] << " " << m_sysdata->m_disp[3 * i + 1] << " " << m_sysdata->m_disp[3 * i + 2] << std::endl;
}

// ------------------------------------------------------------------------------------ exports
namespace {

template <ThermoKind K, int D>
struct ThermoT : ThermoObservable {
    ThermoT(std::string n, std::string t, bool local) : ThermoObservable(K, D, std::move(n), std::move(t), local) {}
};

template <ThermoKind K, int D>
void export_thermo_obs(py::module_& m, const char* name)
{
    using T = ThermoT<K, D>;
    py::class_<T, ThermoObservable, std::shared_ptr<T>>(m, name).def(py::init<std::string, std::string, bool>());
}

template <FieldKind K, int D>
struct FieldT : FieldObservable {
    FieldT(std::string n, std::string t, bool local, int gridsize) : FieldObservable(K, D, std::move(n), std::move(t), local, gridsize) {}
};

template <FieldKind K, int D>
void export_field_obs(py::module_& m, const char* name)
{
    using T = FieldT<K, D>;
    py::class_<T, FieldObservable, std::shared_ptr<T>>(m, name).def(py::init<std::string, std::string, bool, int>());
}

}  // namespace

void export_calculators(py::module_& m)
{
    py::class_<Calculator, std::shared_ptr<Calculator>>(m, "Calculator")
        .def(py::init<std::shared_ptr<ParticleSystem>, std::shared_ptr<PairPotential>>())
        .def("setSystemData", &Calculator::setSystemData)
        .def("compute", &Calculator::compute)
        .def_readonly("max_rcut", &Calculator::max_rcut);

    py::class_<ThermoObservable, Observable, std::shared_ptr<ThermoObservable>>(m, "ThermoObservable")
        .def("save", &ThermoObservable::save)
        .def("clear", &ThermoObservable::clear)
        .def("values", &ThermoObservable::values);
    export_thermo_obs<ThermoKind::Energy, 2>(m, "GlobalPotentialEnergy2D");
    export_thermo_obs<ThermoKind::Energy, 3>(m, "GlobalPotentialEnergy3D");
    export_thermo_obs<ThermoKind::Virial, 2>(m, "GlobalVirialStress2D");
    export_thermo_obs<ThermoKind::Virial, 3>(m, "GlobalVirialStress3D");
    export_thermo_obs<ThermoKind::ElasticVirial, 2>(m, "GlobalElasticVirialStress2D");
    export_thermo_obs<ThermoKind::ElasticVirial, 3>(m, "GlobalElasticVirialStress3D");
    export_thermo_obs<ThermoKind::Kinetic, 2>(m, "GlobalKineticStress2D");
    export_thermo_obs<ThermoKind::Kinetic, 3>(m, "GlobalKineticStress3D");
    export_thermo_obs<ThermoKind::Born, 2>(m, "GlobalBornTensor2D");
    export_thermo_obs<ThermoKind::Born, 3>(m, "GlobalBornTensor3D");

    py::class_<ThermoCalculator, Calculator, std::shared_ptr<ThermoCalculator>>(m, "ThermoCalculator")
        .def(py::init<std::shared_ptr<ParticleSystem>, std::shared_ptr<PairPotential>>())
        .def("compute", &ThermoCalculator::compute)
        .def("setSystemData", &ThermoCalculator::setSystemData)
        .def("addObservable", &ThermoCalculator::addObservable);

    py::class_<FieldObservable, Observable, std::shared_ptr<FieldObservable>>(m, "FieldObservable")
        .def("addGridpoints", &FieldObservable::addGridpoints)
        .def("gettostring", &FieldObservable::gettostring)
        .def("save", &FieldObservable::save)
        .def("clear", &FieldObservable::clear_point)
        .def("values", &FieldObservable::values)
        // additive extension: load values computed elsewhere (re-formatting a saved binary dump; the formatter's tests)
        .def("set_values", [](FieldObservable& f, py::array_t<double, py::array::c_style | py::array::forcecast> a) {
            if ((std::size_t)a.size() != f.npoints * (std::size_t)f.ncomp) throw std::invalid_argument("[ERROR] set_values: expected npoints x ncomp values");
            f.val.assign(a.data(), a.data() + a.size());
        });
    export_field_obs<FieldKind::Virial, 2>(m, "VirialStressField2D");
    export_field_obs<FieldKind::Virial, 3>(m, "VirialStressField3D");
    export_field_obs<FieldKind::ElasticVirial, 2>(m, "ElasticVirialStressField2D");
    export_field_obs<FieldKind::ElasticVirial, 3>(m, "ElasticVirialStressField3D");
    export_field_obs<FieldKind::Kinetic, 2>(m, "KineticStressField2D");
    export_field_obs<FieldKind::Kinetic, 3>(m, "KineticStressField3D");

    py::class_<IrvingKirkwood, Calculator, std::shared_ptr<IrvingKirkwood>>(m, "IrvingKirkwood")
        .def(py::init<std::shared_ptr<ParticleSystem>, std::shared_ptr<PairPotential>, std::shared_ptr<CoarseGrainFunction>,
                      std::shared_ptr<Communicator>>())
        .def("compute", &IrvingKirkwood::compute_points)
        .def("compute_cartesian", &IrvingKirkwood::compute_cartesian)
        .def("setSystemData", &IrvingKirkwood::setSystemData)
        .def("addObservable", &IrvingKirkwood::addObservable)
        .def("printDisplacement", &IrvingKirkwood::printDisplacement);
}

}  // namespace pgt_host
