// module.cpp -- the single pybind11 extension behind pyglasstools' five private modules
// (_pyglasstools, _potential, _cgfunc, _thermo, _irvingkirkwood are thin Python re-exports of this one,
// so that pybind type information is shared without the reference's RTLD_GLOBAL trick, __init__.py:10-11).
#include "host_common.h"

PYBIND11_MODULE(_pgt_host, m)
{
    m.doc() = "pyglasstools_b200 host layer: reference class surface above the libpgt_cuda.so C-ABI";
    pgt_host::export_core(m);
    pgt_host::export_models(m);
    pgt_host::export_calculators(m);
    pgt_host::export_format(m);
}
