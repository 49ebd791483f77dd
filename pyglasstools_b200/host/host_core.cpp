// host_core.cpp -- device session, SimBox, ParticleSystem, Manager, Communicator / LogFile shims.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <sstream>
#include <thread>

#include "host_common.h"
#include "host_io.h"

namespace pgt_host {

// ------------------------------------------------------------------------------------ device
DeviceSession& DeviceSession::get()
{
    static DeviceSession s;
    return s;
}

pgt_ctx* DeviceSession::ctx()
{
    if (m_ctx) return m_ctx;
    int device = 0;
    if (const char* e = std::getenv("PGT_DEVICE")) device = std::atoi(e);
    else if (const char* e2 = std::getenv("LOCAL_RANK")) device = std::atoi(e2);
    int ndev = 0;
    if (pgt_device_count(&ndev) == PGT_OK && ndev > 0) device %= ndev;
    pgt_ctx* c = nullptr;
    const int rc = pgt_ctx_create(device, &c);
    if (rc != PGT_OK || !c)
        throw std::runtime_error(std::string("[ERROR] pyglasstools_b200: cannot open CUDA device (there is no CPU path): ") +
                                 pgt_last_error(nullptr));
    m_ctx = c;
    return m_ctx;
}

void DeviceSession::check(int rc, const char* what)
{
    if (rc == PGT_OK) return;
    std::string msg = std::string("[ERROR] ") + what + ": " + (m_ctx ? pgt_last_error(m_ctx) : pgt_last_error(nullptr));
    if (rc == PGT_ERR_ARG) throw std::invalid_argument(msg);
    if (rc == PGT_ERR_DATA || rc == PGT_ERR_UNSUPPORTED) throw std::domain_error(msg);
    throw std::runtime_error(msg);
}

void DeviceSession::close()
{
    if (m_ctx) pgt_ctx_destroy(m_ctx);
    m_ctx = nullptr;
    frame_token = bonds_token = 0;
    pot_signature.clear();
}

DeviceSession::~DeviceSession() { /* the CUDA runtime may already be gone at static destruction: leak the ctx */ }

// ------------------------------------------------------------------------------------ SimBox
SimBox::SimBox(double L, const Vec3& o, int d) : dim(d), origin(o), boxsize{ L, L, L } { finish(); }
SimBox::SimBox(const Vec3& b, const Vec3& o, int d) : dim(d), origin(o), boxsize(b) { finish(); }

void SimBox::finish()
{
    // the bounds come from the box as given; only afterwards is the third axis flattened (SimBox.cc:10-18, 48-52)
    for (int k = 0; k < 3; ++k) {
        upperbound[k] = 0.5 * boxsize[k] - origin[k];
        lowerbound[k] = (-1.0) * upperbound[k] - 2.0 * origin[k];
    }
    periodic = { 1, 1, 1 };
    if (dim == 2) {
        boxsize[2] = 1.0;
        periodic[2] = 0;
        origin[2] = 0.0;
        vol = boxsize[0] * boxsize[1];
    } else {
        vol = boxsize[0] * boxsize[1] * boxsize[2];
    }
}

Vec3 SimBox::minImage(const Vec3& v) const
{
    // SimBox.h:55-66 (single-precision rint, as the reference has it)
    Vec3 w = v;
    for (int k = 0; k < 3; ++k)
        if (periodic[k]) w[k] -= boxsize[k] * rintf((float)(w[k] / boxsize[k]));
    return w;
}

// ------------------------------------------------------------------------------------ ParticleSystem
static std::atomic<std::uint64_t> g_next_uid{ 1 };

static void copy_rows3(const DArray& a, std::size_t n, std::vector<double>& out, const char* what)
{
    if (a.ndim() != 2 || (std::size_t)a.shape(0) != n || a.shape(1) < 3)
        throw std::invalid_argument(std::string("[ERROR]: Size of ") + what + " array mismatch with # of particles!");
    out.resize(3 * n);
    auto r = a.unchecked<2>();
    for (std::size_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) out[3 * i + k] = r(i, k);
}

static void copy_vec(const DArray& a, std::size_t n, std::vector<double>& out, const char* what)
{
    if (a.ndim() != 1 || (std::size_t)a.shape(0) != n)
        throw std::invalid_argument(std::string("[ERROR]: Size of ") + what + " array mismatch with # of particles!");
    out.assign(a.data(), a.data() + n);
}

ParticleSystem::ParticleSystem(std::shared_ptr<SimBox> box, unsigned int n, py::array diameter, py::array mass, py::array position,
                               py::array velocity)
    : simbox(std::move(box)), m_uid(g_next_uid++)
{
    if (!simbox) throw std::invalid_argument("[ERROR]: ParticleSystem needs a SimBox");
    assign(n, std::move(diameter), std::move(mass), std::move(position), std::move(velocity));
}

namespace {
// A frame array as the copy loop reads it: float32 or float64 rows in C order, read in place (GSD chunks are float32: no
// float64 temporary of the whole frame); anything else goes through numpy's cast once.
struct FrameArray {
    py::array keep;
    const float* f = nullptr;
    const double* d = nullptr;
    std::size_t row = 1;                                        // elements per row
    double at(std::size_t i) const { return f ? (double)f[i] : d[i]; }
};

FrameArray frame_array(py::array a, std::size_t n, int ndim, const char* what)
{
    if (a.ndim() != ndim || (std::size_t)a.shape(0) != n || (ndim == 2 && a.shape(1) < 3))
        throw std::invalid_argument(std::string("[ERROR]: Size of ") + what + " array mismatch with # of particles!");
    FrameArray r;
    const bool c_order = (a.flags() & py::array::c_style) != 0;
    if (c_order && py::isinstance<py::array_t<float>>(a)) {
        r.keep = a;
        r.f = static_cast<const float*>(a.data());
    } else {
        r.keep = DArray(a);                                     // forcecast: one float64 copy
        r.d = static_cast<const double*>(r.keep.data());
    }
    r.row = ndim == 2 ? (std::size_t)r.keep.shape(1) : 1;
    return r;
}

template <class F>
void parallel_chunks(std::size_t n, F&& f)
{
    const unsigned hw = std::thread::hardware_concurrency();
    const int nt = (int)std::max<std::size_t>(1, std::min<std::size_t>({ hw ? hw : 1, 16, n / 16384 + 1 }));
    if (nt <= 1) { f((std::size_t)0, n, 0); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back([&f, n, nt, t] { f(n * t / nt, n * (t + 1) / nt, t); });
    f((std::size_t)0, n / nt, 0);
    for (auto& x : th) x.join();
}
}  // namespace

// One pass over the frame, on several host threads for large frames: promote to FP64, Aboria's enforce_domain
// (NeighbourSearchBase.h:208-237: periodic axes wrapped into [low, high) by repeated +-(high - low); a particle with a
// non-finite coordinate or outside a non-periodic axis is deleted), velocity / mass flags, zero displacements.
void ParticleSystem::assign(unsigned int n_, py::array diameter, py::array mass, py::array position, py::array velocity)
{
    const std::size_t n = n_;
    const FrameArray D = frame_array(std::move(diameter), n, 1, "diameter"), M = frame_array(std::move(mass), n, 1, "mass");
    const FrameArray P = frame_array(std::move(position), n, 2, "position"), V = frame_array(std::move(velocity), n, 2, "velocity");
    if (m_diam.capacity() < n || m_pos.capacity() < 3 * n) unpin();          // the vectors are about to move
    m_diam.resize(n); m_mass.resize(n);
    m_pos.resize(3 * n); m_vel.resize(3 * n); m_disp.resize(3 * n);
    const SimBox& b = *simbox;
    std::vector<unsigned char> dead_any(64, 0), vel_any(64, 0), mass_any(64, 0);
    std::vector<unsigned char> alive(n, 1);
    {
        py::gil_scoped_release nogil;
        parallel_chunks(n, [&](std::size_t i0, std::size_t i1, int t) {
            bool anyv = false, anym = false, anydead = false;
            for (std::size_t i = i0; i < i1; ++i) {
                bool ok = true;
                double r[3];
                for (int k = 0; k < 3; ++k) {
                    r[k] = P.at(P.row * i + k);
                    const double lo = b.lowerbound[k], hi = b.upperbound[k], L = hi - lo;
                    if (!std::isfinite(r[k])) { ok = false; continue; }
                    if (b.periodic[k]) {
                        int guard = 0;
                        while (r[k] < lo && ++guard < (1 << 20)) r[k] += L;
                        while (r[k] >= hi && ++guard < (1 << 20)) r[k] -= L;
                    } else if (r[k] < lo || r[k] >= hi) ok = false;
                }
                for (int k = 0; k < 3; ++k) {
                    const double v = V.at(V.row * i + k);
                    m_pos[3 * i + k] = r[k];
                    m_vel[3 * i + k] = v;
                    m_disp[3 * i + k] = 0.0;
                    anyv |= v != 0.0;
                }
                const double m = M.at(i);
                m_diam[i] = D.at(i);
                m_mass[i] = m;
                anym |= m != 1.0;
                if (!ok) { alive[i] = 0; anydead = true; }
            }
            dead_any[(std::size_t)t] = anydead; vel_any[(std::size_t)t] = anyv; mass_any[(std::size_t)t] = anym;
        });
    }
    bool some_dead = false;
    for (unsigned char c : dead_any) some_dead |= c != 0;
    if (some_dead) {                                            // rare: compact in place, order kept
        std::size_t w = 0;
        for (std::size_t i = 0; i < n; ++i) {
            if (!alive[i]) continue;
            for (int k = 0; k < 3; ++k) {
                m_pos[3 * w + k] = m_pos[3 * i + k];
                m_vel[3 * w + k] = m_vel[3 * i + k];
            }
            m_diam[w] = m_diam[i];
            m_mass[w] = m_mass[i];
            ++w;
        }
        m_pos.resize(3 * w); m_vel.resize(3 * w); m_disp.resize(3 * w);
        m_diam.resize(w); m_mass.resize(w);
        refresh_flags();                                        // the deleted particles' velocities and masses do not count
    } else {
        has_velocity = false;
        unit_mass = true;
        for (unsigned char c : vel_any) has_velocity |= c != 0;
        for (unsigned char c : mass_any) unit_mass &= c == 0;
    }
    touch();
}

void ParticleSystem::refresh_flags()
{
    has_velocity = false;
    for (double v : m_vel)
        if (v != 0.0) { has_velocity = true; break; }
    unit_mass = true;
    for (double m : m_mass)
        if (m != 1.0) { unit_mass = false; break; }
}

// Aboria's enforce_domain (NeighbourSearchBase.h:208-237): periodic axes are wrapped into [low, high) by
// repeated +-(high-low); a particle with a non-finite coordinate, or outside a non-periodic axis, is deleted.
void ParticleSystem::enforce_domain()
{
    const SimBox& b = *simbox;
    const std::size_t n = m_diam.size();
    std::size_t w = 0;
    for (std::size_t i = 0; i < n; ++i) {
        bool alive = true;
        double r[3];
        for (int k = 0; k < 3; ++k) {
            r[k] = m_pos[3 * i + k];
            const double lo = b.lowerbound[k], hi = b.upperbound[k], L = hi - lo;
            if (!std::isfinite(r[k])) { alive = false; break; }
            if (b.periodic[k]) {
                int guard = 0;
                while (r[k] < lo && ++guard < (1 << 20)) r[k] += L;
                while (r[k] >= hi && ++guard < (1 << 20)) r[k] -= L;
            } else if (r[k] < lo || r[k] >= hi) {
                alive = false;
                break;
            }
        }
        if (!alive) continue;
        for (int k = 0; k < 3; ++k) {
            m_pos[3 * w + k] = r[k];
            m_vel[3 * w + k] = m_vel[3 * i + k];
            m_disp[3 * w + k] = m_disp[3 * i + k];
        }
        m_diam[w] = m_diam[i];
        m_mass[w] = m_mass[i];
        ++w;
    }
    m_pos.resize(3 * w); m_vel.resize(3 * w); m_disp.resize(3 * w);
    m_diam.resize(w); m_mass.resize(w);
}

static py::array_t<double> rows3_to_numpy(const std::vector<double>& v)
{
    py::array_t<double> a({ (py::ssize_t)(v.size() / 3), (py::ssize_t)3 });
    if (!v.empty()) std::memcpy(a.mutable_data(), v.data(), v.size() * sizeof(double));
    return a;
}

void ParticleSystem::setMass(const std::vector<double>& v)
{
    if (v.size() != size()) throw std::invalid_argument("[ERROR]: Size of mass array mismatch with # of particles!");
    m_mass = v;
    refresh_flags();
    touch();
}

void ParticleSystem::setDiameter(const std::vector<double>& v)
{
    if (v.size() != size()) throw std::invalid_argument("[ERROR]: Size of diameter array mismatch with # of particles!");
    if (m_diam.capacity() < v.size()) unpin();
    m_diam = v;
    touch();
}

py::array_t<double> ParticleSystem::getDisplacement() const { return rows3_to_numpy(m_disp); }
void ParticleSystem::setDisplacement(DArray v) { copy_rows3(v, size(), m_disp, "displacement"); }
py::array_t<double> ParticleSystem::getAtomPosition() const { return rows3_to_numpy(m_pos); }
py::array_t<double> ParticleSystem::getAtomVelocity() const { return rows3_to_numpy(m_vel); }

void ParticleSystem::setAtomPosition(DArray v)
{
    copy_rows3(v, size(), m_pos, "position");
    touch();   // like the reference, the search structure is only refreshed by moveParticles/update; we re-wrap on upload
}

void ParticleSystem::setAtomVelocity(DArray v)
{
    copy_rows3(v, size(), m_vel, "velocity");
    refresh_flags();
    touch();
}

void ParticleSystem::moveParticles(DArray delta)
{
    std::vector<double> d;
    copy_rows3(delta, size(), d, "position");
    for (std::size_t k = 0; k < m_pos.size(); ++k) m_pos[k] += d[k];
    enforce_domain();
    touch();
}

void ParticleSystem::displaceParticles()
{
    for (std::size_t k = 0; k < m_pos.size(); ++k) m_pos[k] += m_disp[k];
    enforce_domain();
    touch();
}

// Utility getters (not on the pair-sum path): Aboria euclidean_search semantics, Search.h:186-193, 438-447 --
// every periodic image k in {-1,0,1}, dx = p - (c + k L), inclusive radius test.
template <class F>
static void for_each_in_radius(const ParticleSystem& ps, const double c[3], double radius, F&& f)
{
    const SimBox& b = *ps.simbox;
    const double r2 = radius * radius;
    int k0[3], k1[3];
    for (int a = 0; a < 3; ++a) { k0[a] = b.periodic[a] ? -1 : 0; k1[a] = b.periodic[a] ? 1 : 0; }
    const std::size_t n = ps.size();
    for (int kx = k0[0]; kx <= k1[0]; ++kx)
        for (int ky = k0[1]; ky <= k1[1]; ++ky)
            for (int kz = k0[2]; kz <= k1[2]; ++kz) {
                const double cc[3] = { c[0] + kx * (b.upperbound[0] - b.lowerbound[0]), c[1] + ky * (b.upperbound[1] - b.lowerbound[1]),
                                       c[2] + kz * (b.upperbound[2] - b.lowerbound[2]) };
                for (std::size_t i = 0; i < n; ++i) {
                    const double dx = ps.m_pos[3 * i] - cc[0], dy = ps.m_pos[3 * i + 1] - cc[1], dz = ps.m_pos[3 * i + 2] - cc[2];
                    if (dx * dx + dy * dy + dz * dz <= r2) f(i, dx, dy, dz);
                }
            }
}

std::vector<unsigned int> ParticleSystem::getNeighbors(const std::vector<double>& point, double radius) const
{
    if (point.size() < 2) throw std::invalid_argument("[ERROR]: point needs at least two coordinates");
    const double c[3] = { point[0], point[1], point.size() > 2 ? point[2] : 0.0 };
    std::vector<unsigned int> ids;
    for_each_in_radius(*this, c, radius, [&](std::size_t i, double, double, double) { ids.push_back((unsigned int)i); });
    return ids;
}

std::vector<Vec3> ParticleSystem::getNeighborsDistance(int tag, double radius) const
{
    if (tag < 0 || (std::size_t)tag >= size()) throw std::invalid_argument("[ERROR]: particle tag out of range");
    const double c[3] = { m_pos[3 * tag], m_pos[3 * tag + 1], m_pos[3 * tag + 2] };
    std::vector<Vec3> out;
    for_each_in_radius(*this, c, radius, [&](std::size_t i, double dx, double dy, double dz) {
        if ((int)i != tag) out.push_back(Vec3{ -dx, -dy, -dz });      // ParticleSystem.cc:180-186
    });
    return out;
}

// ------------------------------------------------------------------------------------ Manager
// argv "-name value" scanning and notice levels (Manager.h:19-104, 127-201); single process -> rank 0.
class Manager {
public:
    Manager()
    {
        py::list argv = py::module_::import("sys").attr("argv");
        for (std::size_t i = 0; i + 1 < (std::size_t)py::len(argv); ++i) {
            const std::string key = py::str(argv[i]);
            if (key == "-notice_level") {
                const std::string val = py::str(argv[i + 1]);
                char* end = nullptr;
                const long v = std::strtol(val.c_str(), &end, 10);
                if (end == val.c_str() || *end != '\0' || v < 0) throw std::domain_error("[ERROR] Invalid value for notice_level");
                m_notice_level = (unsigned int)v;
            }
        }
    }
    void setNoticeLevel(int l) { m_notice_level = (unsigned int)l; }
    unsigned int getNoticeLevel() const { return m_notice_level; }
    void py_notice(unsigned int level, const std::string& message) const
    {
        if (level <= m_notice_level && level > 1) py::print("notice(" + std::to_string(level) + "): " + message);
    }

private:
    unsigned int m_notice_level = 1;
};

// ------------------------------------------------------------------------------------ Communicator / LogFile
// MPIInterface.h:108-431 reduced to what one process per GPU needs.  Under torchrun the rank / size come from
// RANK / WORLD_SIZE; the data-path collectives are NCCL calls inside libpgt_cuda.so, not Python-level messages.
Communicator::Communicator()
{
    if (const char* r = std::getenv("RANK")) m_rank = std::atoi(r);
    if (const char* s = std::getenv("WORLD_SIZE")) m_size = std::max(1, std::atoi(s));
    if (m_rank < 0 || m_rank >= m_size) { m_rank = 0; m_size = 1; }
}

void Communicator::barrier() const
{
    DeviceSession& s = DeviceSession::get();
    // nothing to wait for across processes at this level; make the device quiescent so that files are consistent
    try { pgt_synchronize(s.ctx()); } catch (const std::exception&) {}
}

void Communicator::abort(int code) const
{
    std::fflush(stdout);
    std::fflush(stderr);
    std::_Exit(code);
}

LogFile::LogFile(std::shared_ptr<Communicator> comm, const std::string& filename) : m_comm(std::move(comm))
{
    m_file = std::fopen(filename.c_str(), "ab");        // MPI_MODE_CREATE | WRONLY | APPEND (MPIFile.h:35)
    if (!m_file) throw std::runtime_error("[ERROR] cannot open log file " + filename);
    std::setvbuf(m_file, nullptr, _IOFBF, 1 << 20);
}

LogFile::~LogFile()
{
    if (m_file) std::fclose(m_file);
}

void LogFile::write_shared(const std::string& s)
{
    if (!s.empty()) std::fwrite(s.data(), 1, s.size(), m_file);
}

void LogFile::flush() { std::fflush(m_file); }

// ------------------------------------------------------------------------------------ exports
void export_core(py::module_& m)
{
    py::class_<SimBox, std::shared_ptr<SimBox>>(m, "SimBox")
        .def(py::init<double, Vec3, int>())
        .def(py::init<Vec3, Vec3, int>())
        .def("minImage", &SimBox::minImage)
        .def_readwrite("dim", &SimBox::dim)
        .def_readwrite("origin", &SimBox::origin)
        .def_readwrite("boxsize", &SimBox::boxsize)
        .def_readwrite("upperbound", &SimBox::upperbound)
        .def_readwrite("lowerbound", &SimBox::lowerbound)
        .def_readwrite("periodic", &SimBox::periodic)
        .def_readwrite("vol", &SimBox::vol);

    py::class_<ParticleSystem, std::shared_ptr<ParticleSystem>>(m, "ParticleSystem")
        // numpy arrays are read in place (float32 GSD chunks without a float64 temporary); lists and other sequences go
        // through numpy's conversion, as the reference's std::vector arguments accept them (ParticleSystem.cc:191-205)
        .def(py::init<std::shared_ptr<SimBox>, unsigned int, py::array, py::array, py::array, py::array>())
        .def(py::init([](std::shared_ptr<SimBox> box, unsigned int n, DArray d, DArray m, DArray p, DArray v) {
            return std::make_shared<ParticleSystem>(std::move(box), n, py::array(d), py::array(m), py::array(p), py::array(v));
        }))
        .def("getMass", &ParticleSystem::getMass)
        .def("setMass", &ParticleSystem::setMass)
        .def("getDiameter", &ParticleSystem::getDiameter)
        .def("setDiameter", &ParticleSystem::setDiameter)
        .def("getDisplacement", &ParticleSystem::getDisplacement)
        .def("setDisplacement", &ParticleSystem::setDisplacement)
        .def("setAtomPosition", &ParticleSystem::setAtomPosition)
        .def("setAtomVelocity", &ParticleSystem::setAtomVelocity)
        .def("getAtomPosition", &ParticleSystem::getAtomPosition)
        .def("getAtomVelocity", &ParticleSystem::getAtomVelocity)
        .def("moveParticles", &ParticleSystem::moveParticles)
        .def("moveParticles", &ParticleSystem::displaceParticles)
        .def("getNeighbors", &ParticleSystem::getNeighbors)
        .def("getNeighborsDistance", &ParticleSystem::getNeighborsDistance)
        .def("update", &ParticleSystem::assign)
        .def("update", [](ParticleSystem& ps, unsigned int n, DArray d, DArray m, DArray p, DArray v) {
            ps.assign(n, py::array(d), py::array(m), py::array(p), py::array(v));
        })
        .def("__len__", &ParticleSystem::size);

    py::class_<Observable, std::shared_ptr<Observable>>(m, "Observable")
        .def(py::init<std::string, std::string, bool, int>())
        .def_readwrite("name", &Observable::name)
        .def_readwrite("type", &Observable::type)
        .def_readwrite("islocal", &Observable::islocal)
        .def_readwrite("dim", &Observable::dim);

    py::class_<Manager, std::shared_ptr<Manager>>(m, "Manager")
        .def(py::init<>())
        .def("notice", &Manager::py_notice)
        .def("setNoticeLevel", &Manager::setNoticeLevel)
        .def("getNoticeLevel", &Manager::getNoticeLevel);

    py::class_<Communicator, std::shared_ptr<Communicator>>(m, "Communicator")
        .def(py::init<>())
        .def("splitPartitions", [](Communicator&, unsigned int) {})
        .def("getPartition", [](const Communicator&) { return 0u; })
        .def("getNRanks", &Communicator::size)
        .def("getRank", &Communicator::rank)
        .def("getSizeGlobal", &Communicator::size)
        .def("getRankGlobal", &Communicator::rank)
        .def("barrier", &Communicator::barrier)
        .def("abort", &Communicator::abort)
        // scatter_v(list_of_chunks, root): this process keeps its own chunk (irvingkirkwood/__init__.py:59)
        .def("scatter_v", [](const Communicator& c, py::sequence chunks, unsigned int) -> py::object {
            if ((int)py::len(chunks) != c.size()) throw std::runtime_error("[ERROR] scatter_v: one chunk per rank expected");
            return chunks[c.rank()];
        })
        .def("all_gather_v", [](const Communicator& c, py::object mine) {
            if (c.size() != 1) throw std::runtime_error("[ERROR] all_gather_v across processes is done on device (pgt_comm_allgather_field)");
            py::list l; l.append(mine); return l;
        })
        .def("bcast", [](const Communicator&, py::object v, unsigned int) { return v; })
        .def("send", [](const Communicator& c, bool, unsigned int, unsigned int) {
            if (c.size() != 1) throw std::runtime_error("[ERROR] point-to-point messages are not part of the one-process-per-GPU design");
        })
        .def("recv", [](const Communicator& c, unsigned int, unsigned int) -> bool {
            if (c.size() != 1) throw std::runtime_error("[ERROR] point-to-point messages are not part of the one-process-per-GPU design");
            return true;
        });

    py::class_<LogFile, std::shared_ptr<LogFile>>(m, "MPILogFile")
        .def(py::init<std::shared_ptr<Communicator>, std::string>())
        .def("write_shared", &LogFile::write_shared)
        .def("flush", &LogFile::flush);

    m.def("device_close", []() { DeviceSession::get().close(); }, "release the CUDA context of this process");
    // bytewise equality of two C-contiguous float64 arrays on several host threads (ikcalculator.run checks every frame that
    // its G x 3 lattice is still the one it generated; np.array_equal costs 5 ms per 1 M points)
    m.def("array_equal_fast", [](DArray a, DArray b) {
        if (a.ndim() != b.ndim() || a.size() != b.size()) return false;
        for (py::ssize_t k = 0; k < a.ndim(); ++k)
            if (a.shape(k) != b.shape(k)) return false;
        const char* pa = reinterpret_cast<const char*>(a.data());
        const char* pb = reinterpret_cast<const char*>(b.data());
        const std::size_t bytes = (std::size_t)a.size() * sizeof(double);
        std::atomic<int> differ{ 0 };
        if (bytes < ((std::size_t)4 << 20)) return std::memcmp(pa, pb, bytes) == 0;
        py::gil_scoped_release nogil;
        const std::size_t units = bytes >> 6;                         // 64-byte units: a thread per 4 MB
        parallel_chunks(units, [&](std::size_t i0, std::size_t i1, int) {
            if (std::memcmp(pa + (i0 << 6), pb + (i0 << 6), (i1 - i0) << 6) != 0) differ.store(1);
        });
        if (std::memcmp(pa + (units << 6), pb + (units << 6), bytes - (units << 6)) != 0) differ.store(1);
        return differ.load() == 0;
    });
    m.def("device_launches", []() { return (long long)pgt_kernel_launches(DeviceSession::get().ctx()); });
}

}  // namespace pgt_host
