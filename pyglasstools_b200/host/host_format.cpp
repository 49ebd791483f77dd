// host_format.cpp -- text formatting of field dumps (io.fieldlogger.save_perrank, io.py:112-128) done in
// one multithreaded C++ pass per frame instead of 3 G Python->C++ calls: byte-identical lines
//   "1 {x} {y} [{z} ]" + for each requested component  ostream<<value<<" "  + "\n"
// where {x} is Python's str(float) (shortest round-trip repr) and the values use the C++ ostream default
// (%g with 6 significant digits), exactly as CoarseGrainedField.h:160-183 prints them.
#include <charconv>
#include <cmath>
#include <cstdio>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include <algorithm>

#include "host_common.h"
#include "host_io.h"

namespace pgt_host {

std::string ostream_double(double v)
{
    char buf[40];
    const int n = std::snprintf(buf, sizeof buf, "%g", v);
    return std::string(buf, (std::size_t)n);
}

static inline char* put_g_chars(char* p, char* end, double v);
// the bulk formatter's route for one value (tests compare it with ostream_double)
std::string ostream_double_fast(double v)
{
    char buf[48];
    return std::string(buf, (std::size_t)(put_g_chars(buf, buf + sizeof buf, v) - buf));
}

// Python's float.__repr__: shortest digits that round-trip; scientific when the decimal exponent is < -4 or >= 16.
std::string repr_double(double v)
{
    if (std::isnan(v)) return "nan";
    if (std::isinf(v)) return v < 0 ? "-inf" : "inf";
    char sci[40];
    auto r = std::to_chars(sci, sci + sizeof sci, v, std::chars_format::scientific);   // d[.ddd]e[+-]XX, shortest
    std::string s(sci, r.ptr);
    const std::size_t epos = s.find('e');
    const int exp10 = std::atoi(s.c_str() + epos + 1);
    if (exp10 < -4 || exp10 >= 16) {
        // Python prints the mantissa as is and at least two exponent digits, like to_chars
        return s;
    }
    char fix[64];
    auto rf = std::to_chars(fix, fix + sizeof fix, v, std::chars_format::fixed);       // shortest fixed notation
    std::string f(fix, rf.ptr);
    if (f.find('.') == std::string::npos) f += ".0";
    return f;
}

// printf("%g") characters of one value.  Fast path: six significant digits from ONE multiplication by an exactly
// representable power of ten and a round-to-nearest, taken only when the scaled value is not within 1e-6 of a rounding
// tie (there the exact decimal expansion decides: std::to_chars(general, 6), specified to produce printf's "%.6g").
static inline char* put_g_chars(char* p, char* end, double v)
{
    static const double P10[23] = { 1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                                    1e17, 1e18, 1e19, 1e20, 1e21, 1e22 };
    const double a = std::fabs(v);
    if (a >= 1e-15 && a < 1e20) {
        int e2;
        std::frexp(a, &e2);
        int e10 = (int)std::floor((e2 - 1) * 0.30102999566398120);           // floor(log10 a) or one less
        for (int tries = 0; tries < 2; ++tries) {
            const int sh = 5 - e10;                                           // a * 10^sh in [1e5, 1e6)
            if (sh < -22 || sh > 22) break;
            const double scaled = sh >= 0 ? a * P10[sh] : a / P10[-sh];
            if (scaled >= 999999.4999) { if (scaled < 999999.5001) break; ++e10; continue; }   // at the tie: exact route
            if (scaled < 99999.5) { if (scaled < 99999.4999) { --e10; continue; } break; }
            const double fl = std::floor(scaled), frac = scaled - fl;
            if (std::fabs(frac - 0.5) < 1e-6) break;                          // near a tie: the slow, exact route
            long long m = (long long)fl + (frac > 0.5 ? 1 : 0);              // 100000 .. 1000000
            if (m >= 1000000) { m = 100000; ++e10; }
            char d[6];
            for (int k = 5; k >= 0; --k) { d[k] = (char)('0' + m % 10); m /= 10; }
            int nd = 6;
            while (nd > 1 && d[nd - 1] == '0') --nd;                          // %g drops trailing zeros
            if (std::signbit(v)) *p++ = '-';
            if (e10 < -4 || e10 >= 6) {                                       // d[.ddddd]e+XX
                *p++ = d[0];
                if (nd > 1) { *p++ = '.'; for (int k = 1; k < nd; ++k) *p++ = d[k]; }
                *p++ = 'e';
                int x = e10;
                if (x < 0) { *p++ = '-'; x = -x; } else *p++ = '+';
                if (x >= 100) { *p++ = (char)('0' + x / 100); x %= 100; }
                *p++ = (char)('0' + x / 10);
                *p++ = (char)('0' + x % 10);
            } else if (e10 >= 0) {                                            // ddd[.ddd]
                for (int k = 0; k <= e10; ++k) *p++ = k < nd ? d[k] : '0';
                if (nd > e10 + 1) { *p++ = '.'; for (int k = e10 + 1; k < nd; ++k) *p++ = d[k]; }
            } else {                                                          // 0.000ddd
                *p++ = '0'; *p++ = '.';
                for (int k = -1; k > e10; --k) *p++ = '0';
                for (int k = 0; k < nd; ++k) *p++ = d[k];
            }
            return p;
        }
    }
    return std::to_chars(p, end, v, std::chars_format::general, 6).ptr;
}

static inline char* put_g(char* p, char* end, double v)
{
    p = put_g_chars(p, end, v);
    *p = ' ';
    return p + 1;
}

// The lines of grid points [g0, g1).
static void format_range(const double* P, std::ptrdiff_t pstride, int ncoord, const std::vector<const double*>& cols,
                         const std::vector<int>& ncomp, const std::vector<int>& comps, std::size_t g0, std::size_t g1, std::string& out)
{
    out.clear();                                             // pooled: keeps the capacity the last frame needed
    if (out.capacity() == 0) out.reserve((g1 - g0) * (6 + 2 * 12 + 12 * cols.size()));
    // coordinates repeat heavily on Cartesian grids (every x ny times in a row, every y once per row): a small
    // direct-mapped cache of their strings, keyed on the bit pattern
    struct Slot { std::uint64_t bits; unsigned char len; char txt[30]; };
    constexpr std::size_t NSLOT = 4096;
    std::vector<Slot> cache(NSLOT, Slot{ ~std::uint64_t(0), 0, { 0 } });     // all-ones is a NaN pattern: formatted uncached
    char buf[64];
    for (std::size_t g = g0; g < g1; ++g) {
        out += "1 ";
        for (int k = 0; k < ncoord; ++k) {
            const double c = P[g * pstride + k];
            std::uint64_t bits;
            std::memcpy(&bits, &c, 8);
            Slot& sl = cache[(std::size_t)((bits * 0x9E3779B97F4A7C15ull) >> 52)];
            if (sl.bits != bits || bits == ~std::uint64_t(0)) {
                const std::string r = repr_double(c);
                if (r.size() + 1 <= sizeof sl.txt) {
                    sl.bits = bits;
                    sl.len = (unsigned char)(r.size() + 1);
                    std::memcpy(sl.txt, r.data(), r.size());
                    sl.txt[r.size()] = ' ';
                } else {
                    out += r;
                    out += ' ';
                    continue;
                }
            }
            out.append(sl.txt, sl.len);
        }
        char* p = buf;
        for (std::size_t f = 0; f < cols.size(); ++f) {
            p = put_g(p, buf + sizeof buf, cols[f][g * ncomp[f] + comps[f]]);
            if (p > buf + 40) { out.append(buf, (std::size_t)(p - buf)); p = buf; }
        }
        *p++ = '\n';
        out.append(buf, (std::size_t)(p - buf));
    }
}

// The frame's lines for points [0, G): the grid is cut into contiguous chunks, one per host thread, and the chunks are
// used in order -- the same bytes as the reference's per-point loop (io.py:112-128).  The chunk strings live in a
// process-wide pool and keep their capacity from frame to frame: a fresh 70 MB of text buffers per frame costs more in
// page faults than the formatting itself.  The pool's mutex is only ever waited for with the GIL released.
namespace {
struct FramePrep {
    py::array_t<double, py::array::c_style | py::array::forcecast> pts;
    const double* P = nullptr;
    std::ptrdiff_t pstride = 3;
    int ncoord = 2;
    std::size_t G = 0;
    std::vector<const double*> cols;
    std::vector<int> ncomp, comps;
    int nt = 1;
};
std::mutex g_parts_mu;
std::vector<std::string> g_parts;
}  // namespace

static FramePrep prepare_frame(DArray points, int dim, const std::vector<std::shared_ptr<FieldObservable>>& fields,
                               const std::vector<int>& comps, int nthreads)
{
    if (points.ndim() != 2 || points.shape(1) < 3) throw std::invalid_argument("[ERROR] points must be (G,3)");
    if (fields.size() != comps.size()) throw std::invalid_argument("[ERROR] one component index per field expected");
    FramePrep f;
    f.G = (std::size_t)points.shape(0);
    for (std::size_t k = 0; k < fields.size(); ++k) {
        if (!fields[k] || fields[k]->npoints != f.G) throw std::invalid_argument("[ERROR] field size does not match the grid");
        if (comps[k] < 0 || comps[k] >= fields[k]->ncomp) throw std::invalid_argument("[ERROR] component index out of range");
        f.cols.push_back(fields[k]->val.data());
        f.ncomp.push_back(fields[k]->ncomp);
    }
    f.comps = comps;
    f.pts = py::array_t<double, py::array::c_style | py::array::forcecast>(points);
    f.P = f.pts.data();
    f.pstride = f.pts.shape(1);
    f.ncoord = dim == 3 ? 3 : 2;
    int nt = nthreads > 0 ? nthreads : (int)std::thread::hardware_concurrency();
    f.nt = std::max(1, std::min<int>({ nt, 64, (int)(f.G / 4096) + 1 }));
    return f;
}

// No GIL needed; the caller holds g_parts_mu.  The grid is cut into four chunks per thread; the threads take chunks in
// order and the CALLING thread consumes each chunk (in order) as soon as it is complete, so that writing the text to the
// file overlaps the formatting of the chunks behind it.
template <class CONSUME>
static void format_chunks(const FramePrep& f, CONSUME&& consume)
{
    const int nt = f.nt;
    const std::size_t G = f.G;
    const int C = nt == 1 ? 1 : 4 * nt;
    if ((int)g_parts.size() < C) g_parts.resize((std::size_t)C);
    if (nt == 1) {
        format_range(f.P, f.pstride, f.ncoord, f.cols, f.ncomp, f.comps, 0, G, g_parts[0]);
        consume(g_parts[0]);
        return;
    }
    std::atomic<int> next{ 0 };
    std::vector<std::atomic<int>> done((std::size_t)C);
    for (auto& d : done) d.store(0, std::memory_order_relaxed);
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t)
        th.emplace_back([&] {
            for (;;) {
                const int c = next.fetch_add(1);
                if (c >= C) break;
                format_range(f.P, f.pstride, f.ncoord, f.cols, f.ncomp, f.comps, G * (std::size_t)c / C, G * (std::size_t)(c + 1) / C, g_parts[(std::size_t)c]);
                done[(std::size_t)c].store(1, std::memory_order_release);
            }
        });
    for (int c = 0; c < C; ++c) {
        while (!done[(std::size_t)c].load(std::memory_order_acquire)) std::this_thread::yield();
        consume(g_parts[(std::size_t)c]);
    }
    for (auto& x : th) x.join();
}

static py::bytes format_field_frame_impl(DArray points, int dim, const std::vector<std::shared_ptr<FieldObservable>>& fields,
                                         const std::vector<int>& comps, int nthreads)
{
    const FramePrep f = prepare_frame(points, dim, fields, comps, nthreads);
    std::string out;
    {
        py::gil_scoped_release nogil;
        std::lock_guard<std::mutex> lk(g_parts_mu);
        format_chunks(f, [&](const std::string& s) { out += s; });
    }
    return py::bytes(out);
}

// The same lines straight into the log file: the chunks are written in order as they are, without being joined or
// handed through Python strings (at G = 1 M that round trip costs more than the formatting).
static void write_field_frame(std::shared_ptr<LogFile> file, DArray points, int dim,
                              const std::vector<std::shared_ptr<FieldObservable>>& fields, const std::vector<int>& comps)
{
    if (!file) throw std::invalid_argument("[ERROR] write_field_frame: no file");
    const FramePrep f = prepare_frame(points, dim, fields, comps, 0);
    py::gil_scoped_release nogil;
    std::lock_guard<std::mutex> lk(g_parts_mu);
    format_chunks(f, [&](const std::string& s) { file->write_shared(s); });
}

static std::string format_field_frame(DArray points, int dim, const std::vector<std::shared_ptr<FieldObservable>>& fields,
                                      const std::vector<int>& comps)
{
    return std::string(format_field_frame_impl(points, dim, fields, comps, 0));
}

void export_format(py::module_& m)
{
    m.def("format_field_frame_ms", [](DArray points, int dim, const std::vector<std::shared_ptr<FieldObservable>>& fields,
                                      const std::vector<int>& comps, int nthreads) {
        // developer hook: milliseconds of the formatting pass alone (no join, no Python bytes, no file)
        const FramePrep f = prepare_frame(points, dim, fields, comps, nthreads);
        py::gil_scoped_release nogil;
        std::lock_guard<std::mutex> lk(g_parts_mu);
        const auto t0 = std::chrono::steady_clock::now();
        format_chunks(f, [](const std::string&) {});
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    });
    m.def("repr_double", &repr_double);
    m.def("ostream_double", &ostream_double);
    m.def("ostream_double_fast", &ostream_double_fast);
    m.def("ostream_double_fast_check", [](py::array_t<double, py::array::c_style | py::array::forcecast> a) {
        // number of values whose fast formatting differs from snprintf("%g") (tests: must be 0)
        long long bad = 0;
        char b1[48], b2[48];
        for (py::ssize_t i = 0; i < a.size(); ++i) {
            const int n1 = std::snprintf(b1, sizeof b1, "%g", a.data()[i]);
            const int n2 = (int)(put_g_chars(b2, b2 + sizeof b2, a.data()[i]) - b2);
            if (n1 != n2 || std::memcmp(b1, b2, (std::size_t)n1)) ++bad;
        }
        return bad;
    });
    m.def("format_field_frame", &format_field_frame, py::arg("points"), py::arg("dim"), py::arg("fields"), py::arg("components"));
    m.def("write_field_frame", &write_field_frame, py::arg("file"), py::arg("points"), py::arg("dim"), py::arg("fields"), py::arg("components"));
    // same lines as bytes, with an explicit thread count (0: all host threads; 1: the serial reference for the tests)
    m.def("format_field_frame_bytes", &format_field_frame_impl, py::arg("points"), py::arg("dim"), py::arg("fields"), py::arg("components"),
          py::arg("nthreads") = 0);
}

}  // namespace pgt_host
