// host_format.cpp -- text formatting of field dumps (io.fieldlogger.save_perrank, io.py:112-128) done in
// one C++ pass per frame instead of 3 G Python->C++ calls: byte-identical lines
//   "1 {x} {y} [{z} ]" + for each requested component  ostream<<value<<" "  + "\n"
// where {x} is Python's str(float) (shortest round-trip repr) and the values use the C++ ostream default
// (%g with 6 significant digits), exactly as CoarseGrainedField.h:160-183 prints them.
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "host_common.h"

namespace pgt_host {

std::string ostream_double(double v)
{
    char buf[40];
    const int n = std::snprintf(buf, sizeof buf, "%g", v);
    return std::string(buf, (std::size_t)n);
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

// Appends the frame's lines for points [0, G) to a single string.
static std::string format_field_frame(DArray points, int dim, const std::vector<std::shared_ptr<FieldObservable>>& fields,
                                      const std::vector<int>& comps)
{
    if (points.ndim() != 2 || points.shape(1) < 3) throw std::invalid_argument("[ERROR] points must be (G,3)");
    if (fields.size() != comps.size()) throw std::invalid_argument("[ERROR] one component index per field expected");
    const std::size_t G = (std::size_t)points.shape(0);
    for (std::size_t f = 0; f < fields.size(); ++f) {
        if (!fields[f] || fields[f]->npoints != G) throw std::invalid_argument("[ERROR] field size does not match the grid");
        if (comps[f] < 0 || comps[f] >= fields[f]->ncomp) throw std::invalid_argument("[ERROR] component index out of range");
    }
    auto P = points.unchecked<2>();
    std::string out;
    out.reserve(G * (24 + 2 * 20 + 14 * fields.size()));
    // coordinates repeat heavily on Cartesian grids: memoise the last strings per axis
    double last[3] = { NAN, NAN, NAN };
    std::string lasts[3];
    char buf[40];
    for (std::size_t g = 0; g < G; ++g) {
        out += "1 ";
        for (int k = 0; k < (dim == 3 ? 3 : 2); ++k) {
            const double c = P(g, k);
            if (!(c == last[k]) || std::signbit(c) != std::signbit(last[k])) { last[k] = c; lasts[k] = repr_double(c); }
            out += lasts[k];
            out += ' ';
        }
        for (std::size_t f = 0; f < fields.size(); ++f) {
            const int n = std::snprintf(buf, sizeof buf, "%g ", fields[f]->val[g * fields[f]->ncomp + comps[f]]);
            out.append(buf, (std::size_t)n);
        }
        out += '\n';
    }
    return out;
}

void export_format(py::module_& m)
{
    m.def("repr_double", &repr_double);
    m.def("ostream_double", &ostream_double);
    m.def("format_field_frame", &format_field_frame, py::arg("points"), py::arg("dim"), py::arg("fields"), py::arg("components"));
}

}  // namespace pgt_host
