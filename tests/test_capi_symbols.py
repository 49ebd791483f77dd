"""The C-ABI library loads and exports every entry point include/pgt_cuda.h declares (no GPU needed).

No compute call is made here; the only device-touching call checked is that context creation
FAILS LOUDLY without a CUDA device (there is no CPU fallback behind the header).
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pgt_cuda.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pgt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_documented_surface():
    names = declared_functions()
    for must in ("pgt_ctx_create", "pgt_frame_upload", "pgt_build_cells", "pgt_build_bonds", "pgt_ik_compute",
                 "pgt_ik_download", "pgt_thermo_compute", "pgt_comm_allgather_field", "pgt_comm_allreduce_sum"):
        assert must in names
    assert len(names) >= 40


def test_library_exports_every_declared_symbol():
    from pyglasstools_b200 import _capi
    lib = C.CDLL(_capi.LIB_PATH)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, "declared in include/pgt_cuda.h but not exported: %s" % missing
    # and the ctypes binding layer binds the same set
    assert sorted(_capi.EXPORTED) == declared_functions()
    _capi.lib()


def test_library_is_built_for_sm_100a_only():
    from pyglasstools_b200 import _capi
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", _capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_scalar_host_getters_match_oracle(orc):
    """pgt_pair_eval / pgt_cg_delta / pgt_cg_bond are plain host functions of the library (the reference's
    getPairEnergy / getPairForceDivR / getDeltaFunc / getBondFunc Python-visible getters)."""
    from pyglasstools_b200 import _capi
    L = _capi.lib()
    rng = np.random.default_rng(5)
    for name, gname in (("poly12", "poly12"), ("poly10", "poly10"), ("lj", "lj"), ("fslj", "forceshifted-lj"),
                        ("polyyukawa", "polyyukawa")):
        P = orc.potential(name)
        prm = np.array(orc.pot_params(P), np.float64)
        n = 200
        di, dj = rng.uniform(0.8, 1.5, n), rng.uniform(0.8, 1.5, n)
        r = rng.uniform(0.5, 1.1, n) * P.scaled_rcut * 0.5 * (di + dj)
        rij = np.stack([r, 0 * r, 0 * r], axis=1)
        want = orc.pair_table(P, rij, di, dj)
        for what, key in enumerate(("energy", "fdivr", "stiff", "rcutsq")):
            got = np.array([L.pgt_pair_eval(_capi.POT_IDS[gname], P.scaled_rcut, prm.ctypes.data_as(C.POINTER(C.c_double)),
                                            len(prm), rij[k].ctypes.data_as(C.POINTER(C.c_double)), di[k], dj[k], what)
                            for k in range(n)])
            assert np.array_equal(got, want[key]), (name, key)
    for kernel in ("octic", "rect", "mollifier"):
        cg = orc.cgfunc(kernel, 2.0, 5)
        d2 = np.linspace(0, 5, 101)
        got = np.array([L.pgt_cg_delta(_capi.CG_IDS[kernel], 2.0, v) for v in d2])
        assert np.array_equal(got, orc.cg_table(cg, d2))
        x = np.array([0.1, -0.2, 0.0]); ri = np.array([0.9, 0.4, 0.0]); rij = np.array([0.7, -0.9, 0.0])
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        got = L.pgt_cg_bond(_capi.CG_IDS[kernel], 2.0, 5, dp(x), dp(ri), dp(rij))
        want = orc.bond(cg, x, ri, rij)
        assert abs(got - want) <= 1e-15 * max(abs(want), 1e-300) + 1e-18


def test_context_creation_fails_loudly_without_a_gpu():
    from pyglasstools_b200 import _capi
    n = C.c_int(0)
    rc = _capi.lib().pgt_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(_capi.PgtError):
        _capi.Context(0)
