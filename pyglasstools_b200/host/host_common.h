// host_common.h -- C++ host layer above the C-ABI (include/pgt_cuda.h).
//
// Re-exposes the reference's pybind11 class names (SimBox, ParticleSystem, PairPotential*, CGFunc*,
// ThermoCalculator, IrvingKirkwood, Global*/...Field* observables, Manager, Communicator, MPILogFile)
// so that pyglasstools' example_scripts run unchanged, while every pair sum is executed by
// libpgt_cuda.so on the GPU.  Nothing in here computes a pair sum on the host: without a CUDA device
// the first compute() throws.
//
// Reference interfaces mirrored (file:line under /root/reference/pyglasstools/):
//   SimBox.cc:3-68, ParticleSystem.cc:3-220, Calculator.h:27-64, Observable.h:12-41, Manager.h:106-215,
//   MPIInterface.cpp:142-163, MPIFile.h:23-66, potential/PairPotential.h:14-152,
//   irvingkirkwood/cgfunc/CoarseGrainFunction.h:14-167, thermo/ThermoCalculator.h:14-100,
//   thermo/ThermoProperty.h:9-225, irvingkirkwood/IrvingKirkwood.h:17-122,
//   irvingkirkwood/CoarseGrainedField.h:8-234
#pragma once
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <array>
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "pgt_cuda.h"

namespace py = pybind11;

namespace pgt_host {

using Vec3 = std::array<double, 3>;
using DArray = py::array_t<double, py::array::c_style | py::array::forcecast>;

// ---------------------------------------------------------------------------------------------
// One pgt_ctx per process (one process per GPU, like one MPI rank per core in the reference).
// Created on first use; the device is PGT_DEVICE, else LOCAL_RANK, else 0.
class DeviceSession {
public:
    static DeviceSession& get();
    pgt_ctx* ctx();                      // throws std::runtime_error when no CUDA device can be opened
    void check(int rc, const char* what);
    // What the device currently holds, so that two calculators sharing a ParticleSystem (thermo + IK)
    // do not upload / rebuild twice.
    std::uint64_t frame_token = 0;       // ParticleSystem uid/version of the uploaded frame + max_rcut
    double frame_max_rcut = 0.0;
    std::string pot_signature;           // model + parameters the half-bond list was built with
    std::uint64_t bonds_token = 0;
    void close();
    ~DeviceSession();

private:
    pgt_ctx* m_ctx = nullptr;
};

// A host allocation page-locked in place (pgt_host_register) for as long as it keeps its address and size: the frame
// arrays of a ParticleSystem that is refilled frame after frame, the holders of the field observables.  Copies from and
// to them then run at link speed instead of through the driver's pageable staging.  Registration failures are not
// errors (the buffer stays pageable).
struct PinnedView {
    void* p = nullptr;
    std::size_t bytes = 0;
    void ensure(const void* ptr, std::size_t n)
    {
        if (ptr == p && n == bytes) return;
        release();
        if (ptr && n >= (std::size_t)1 << 20 && pgt_host_register(const_cast<void*>(ptr), n) == PGT_OK) { p = const_cast<void*>(ptr); bytes = n; }
    }
    void release()
    {
        if (p) pgt_host_unregister(p);
        p = nullptr;
        bytes = 0;
    }
    PinnedView() = default;
    PinnedView(const PinnedView&) = delete;
    PinnedView& operator=(const PinnedView&) = delete;
    ~PinnedView() { release(); }
};

// ---------------------------------------------------------------------------------------------
struct SimBox {
    SimBox(double L, const Vec3& origin, int dim);
    SimBox(const Vec3& boxsize, const Vec3& origin, int dim);
    Vec3 minImage(const Vec3& v) const;
    int dim = 3;
    Vec3 origin{}, boxsize{}, upperbound{}, lowerbound{};
    std::array<int, 3> periodic{ 1, 1, 1 };
    double vol = 0.0;

private:
    void finish();
};

class ParticleSystem {
public:
    ParticleSystem(std::shared_ptr<SimBox> box, unsigned int n, py::array diameter, py::array mass, py::array position, py::array velocity);
    void assign(unsigned int n, py::array diameter, py::array mass, py::array position, py::array velocity);

    std::size_t size() const { return m_diam.size(); }
    std::shared_ptr<SimBox> simbox;
    // host copies (FP64, wrapped into the box exactly as Aboria's init_neighbour_search does)
    std::vector<double> m_pos, m_vel, m_disp;     // 3 per particle
    std::vector<double> m_diam, m_mass;
    bool has_velocity = false;                    // any non-zero velocity component
    bool unit_mass = true;                        // all masses == 1

    mutable PinnedView pin_pos, pin_diam;          // m_pos / m_diam page-locked in place while they keep their address
    void unpin() const { pin_pos.release(); pin_diam.release(); }

    std::uint64_t token() const { return (m_uid << 20) ^ m_version; }
    void touch() { ++m_version; }
    void enforce_domain();                        // NeighbourSearchBase.h:208-237

    // python-facing accessors (ParticleSystem.cc:96-189)
    std::vector<double> getMass() const { return m_mass; }
    void setMass(const std::vector<double>& v);
    std::vector<double> getDiameter() const { return m_diam; }
    void setDiameter(const std::vector<double>& v);
    py::array_t<double> getDisplacement() const;
    void setDisplacement(DArray v);
    py::array_t<double> getAtomPosition() const;
    void setAtomPosition(DArray v);
    py::array_t<double> getAtomVelocity() const;
    void setAtomVelocity(DArray v);
    void moveParticles(DArray delta);
    void displaceParticles();
    std::vector<unsigned int> getNeighbors(const std::vector<double>& point, double radius) const;
    std::vector<Vec3> getNeighborsDistance(int tag, double radius) const;

private:
    std::uint64_t m_uid;
    std::uint64_t m_version = 0;
    void refresh_flags();
};

// ---------------------------------------------------------------------------------------------
struct PairPotential {
    PairPotential(int model, double scaled_rcut, std::vector<double> params)
        : model(model), scaled_rcut(scaled_rcut), params(std::move(params)) {}
    virtual ~PairPotential() = default;
    double eval(const Vec3& rij, double di, double dj, int what) const;
    double getRcut(const Vec3& rij, double di, double dj) const { return eval(rij, di, dj, 3); }
    double getPairForceDivR(const Vec3& rij, double di, double dj) const { return eval(rij, di, dj, 1); }
    double getPairEnergy(const Vec3& rij, double di, double dj) const { return eval(rij, di, dj, 0); }
    double getBondStiffness(const Vec3& rij, double di, double dj) const { return eval(rij, di, dj, 2); }
    void setParams(double val, int num) { params.at(num) = val; }
    std::vector<double> getParams() const { return params; }
    std::string signature() const;
    int model;                // PGT_POT_* ; -1 for the abstract base (every getter returns 0 like the reference)
    double scaled_rcut;
    std::vector<double> params;
};

struct CoarseGrainFunction {
    CoarseGrainFunction(int kernel, int quad_mode, int order, double rcut);
    virtual ~CoarseGrainFunction() = default;
    void setRcut(double r) { cg_rcut = r; }
    double getRcut() const { return cg_rcut; }
    double getDeltaFunc(const Vec3& x, const Vec3& ri) const;
    double getObjFunc(double s, const Vec3& x, const Vec3& ri, const Vec3& rij) const;
    double getBondFunc(const Vec3& x, const Vec3& ri, const Vec3& rij) const;
    int kernel;               // PGT_CG_* ; -1 for the abstract base
    int quad_mode;            // PGT_QUAD_*
    int order;
    double cg_rcut;
    double relerr = 0, abserr = 0;
    int maxiter = 0;
    double norm_scale = 1.0;  // mollifier: the normalisation is frozen at construction (CGFuncMollifier.h:12,17)
    double rcut_at_ctor;
};

// ---------------------------------------------------------------------------------------------
struct Observable {
    Observable() = default;
    Observable(std::string name, std::string type, bool islocal, int dim)
        : name(std::move(name)), type(std::move(type)), islocal(islocal), dim(dim) {}
    virtual ~Observable() = default;
    virtual void clear() {}
    std::string name = "NONE", type = "SCALAR";
    bool islocal = true;
    int dim = 3;
};

class Communicator;
class LogFile;

// thermo holders (ThermoProperty.h): kind selects which slice of the device result is kept
enum class ThermoKind { Energy, Virial, ElasticVirial, Kinetic, Born };
struct ThermoObservable : Observable {
    ThermoObservable(ThermoKind kind, int dim, std::string name, std::string type, bool islocal);
    void clear() override { std::fill(val.begin(), val.end(), 0.0); }
    void save(std::shared_ptr<LogFile> file, int index) const;     // std::to_string -> "%f"
    py::array_t<double> values() const;                            // additive extension: full precision
    unsigned mask() const;
    ThermoKind kind;
    std::vector<double> val;                                       // col-major flat tensor
};

// IK field holders (CoarseGrainedField.h)
enum class FieldKind { Virial, ElasticVirial, Kinetic };
struct FieldObservable : Observable {
    FieldObservable(FieldKind kind, int dim, std::string name, std::string type, bool islocal, int gridsize);
    void addGridpoints(DArray pts);
    void clear_point(int grid_id);
    std::string gettostring(int index, int grid_id) const;         // ostream << double << " "  (6 significant digits)
    void save(std::shared_ptr<LogFile> file, int index, int grid_id) const;
    py::array_t<double> values() const;                            // additive extension: (G, dim*dim) FP64
    FieldKind kind;
    int ncomp;
    std::size_t npoints;
    std::vector<double> val;                                       // npoints x ncomp, col-major per point
    PinnedView pin_val;                                            // val page-locked in place (sized once, at construction)
    std::vector<double> points;                                    // npoints x 3
};

// ---------------------------------------------------------------------------------------------
class Calculator {
public:
    Calculator(std::shared_ptr<ParticleSystem> sys, std::shared_ptr<PairPotential> pot);
    virtual ~Calculator() = default;
    void setSystemData(std::shared_ptr<ParticleSystem> sys) { m_sysdata = std::move(sys); }
    virtual void compute() {}
    double max_rcut;          // frozen at construction (Calculator.h:37-46)

protected:
    // pushes potential + frame to the device and builds the cell list (and the half-bond list on request)
    void prepare_device(bool need_bonds);
    std::shared_ptr<ParticleSystem> m_sysdata;
    std::shared_ptr<PairPotential> m_potential;
};

class ThermoCalculator : public Calculator {
public:
    using Calculator::Calculator;
    void addObservable(std::shared_ptr<ThermoObservable> obs) { m_obs[obs->name] = std::move(obs); }
    void compute() override;

private:
    std::map<std::string, std::shared_ptr<ThermoObservable>> m_obs;
};

class IrvingKirkwood : public Calculator {
public:
    IrvingKirkwood(std::shared_ptr<ParticleSystem> sys, std::shared_ptr<PairPotential> pot,
                   std::shared_ptr<CoarseGrainFunction> cg, std::shared_ptr<Communicator> comm);
    void addObservable(std::shared_ptr<FieldObservable> obs) { m_obs[obs->name] = std::move(obs); }
    void compute_points(DArray gridpoints);
    // additive extension: the i-major Cartesian product xs x ys, restricted to global points [g0,g1)
    void compute_cartesian(DArray xs, DArray ys, long long g0, long long g1);
    void printDisplacement() const;

private:
    void run(long long npoints);
    std::shared_ptr<CoarseGrainFunction> m_cg;
    std::shared_ptr<Communicator> m_comm;
    std::map<std::string, std::shared_ptr<FieldObservable>> m_obs;
};

// ---------------------------------------------------------------------------------------------
void export_core(py::module_& m);        // SimBox, ParticleSystem, Observable, Calculator, Manager, Communicator, MPILogFile
void export_models(py::module_& m);      // PairPotential*, CGFunc*
void export_calculators(py::module_& m); // ThermoCalculator, IrvingKirkwood, Global*, *Field*
void export_format(py::module_& m);      // bulk text formatter for field dumps

std::string repr_double(double v);       // Python's repr(float)
std::string ostream_double(double v);    // C++ ostream default formatting (%g, 6 significant digits)

}  // namespace pgt_host
