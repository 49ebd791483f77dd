"""ctypes bindings for the CPU oracle (oracle/liborc.so) and the reference-math shim.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(pyglasstools_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

MODEL_IDS = {"lj": 0, "fslj": 1, "poly12": 2, "poly18": 3, "poly10": 4, "poly106": 5, "polyyukawa": 6, "polylj": 7}
KERNEL_IDS = {"octic": 0, "rect": 1, "mollifier": 2}


class OrcPot(C.Structure):
    _fields_ = [("model", C.c_int), ("scaled_rcut", C.c_double), ("p", C.c_double * 8)]


class OrcCG(C.Structure):
    _fields_ = [("kernel", C.c_int), ("rcut", C.c_double), ("order", C.c_int), ("mollifier_norm", C.c_double),
                ("adaptive", C.c_int), ("epsabs", C.c_double), ("epsrel", C.c_double), ("limit", C.c_int)]


def build(force=False):
    """Compile liborc.so (and _ref/libpgt_refmath.so when /root/reference exists)."""
    so = os.path.join(_HERE, "liborc.so")
    src = os.path.join(_HERE, "pgt_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liborc.so"], stdout=subprocess.DEVNULL)
    ref = os.path.join(_HERE, "_ref", "libpgt_refmath.so")
    if os.path.isdir("/root/reference/pyglasstools/potential") and (force or not os.path.exists(ref)):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None
_ref = None


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64)) if a is not None else None


def lib():
    """liborc.so (ISO-C reading, -ffp-contract=off: the PARITY build); with PGT_ORACLE_FAST=1 in the environment before the
    first call, liborc_fast.so: the same source built with the reference's own flags (-O3 -march=native,
    CompilerFlagsSetup.cmake:14), the TIMING build of bench.py's CPU legs."""
    global _lib
    if _lib is None:
        build()
        name = "liborc.so"
        if os.environ.get("PGT_ORACLE_FAST") == "1":
            # -march=native binds the build to the CPU it was made on: rebuilt when this host's CPU is another model
            fast = os.path.join(_HERE, "liborc_fast.so")
            src = os.path.join(_HERE, "pgt_oracle.cpp")
            tag = os.path.join(_HERE, "liborc_fast.cpu")
            try:
                cpu = [l for l in open("/proc/cpuinfo") if l.startswith("model name")][0].split(":", 1)[1].strip()
            except Exception:
                cpu = "unknown"
            built_for = open(tag).read().strip() if os.path.exists(tag) else ""
            if not os.path.exists(fast) or os.path.getmtime(fast) < os.path.getmtime(src) or built_for != cpu:
                subprocess.check_call(["make", "-B", "-C", _HERE, "liborc_fast.so"], stdout=subprocess.DEVNULL)
                open(tag, "w").write(cpu + "\n")
            name = "liborc_fast.so"
        L = C.CDLL(os.path.join(_HERE, name))
        L._pgt_build = name
        L.orc_pair_eval.restype = C.c_double
        L.orc_pair_eval.argtypes = [C.POINTER(OrcPot), C.POINTER(C.c_double), C.c_double, C.c_double, C.c_int]
        L.orc_cg_compute.restype = C.c_double
        L.orc_cg_compute.argtypes = [C.POINTER(OrcCG), C.c_double]
        L.orc_delta.restype = C.c_double
        L.orc_bond.restype = C.c_double
        L.orc_system_create.restype = C.c_void_p
        L.orc_system_create.argtypes = [C.c_int, C.POINTER(C.c_double), C.c_int64, C.POINTER(C.c_double),
                                        C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double]
        L.orc_system_destroy.argtypes = [C.c_void_p]
        L.orc_system_n.restype = C.c_int64
        L.orc_system_n.argtypes = [C.c_void_p]
        L.orc_system_vol.restype = C.c_double
        L.orc_system_vol.argtypes = [C.c_void_p]
        L.orc_system_get.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.orc_max_rcut.restype = C.c_double
        L.orc_max_rcut.argtypes = [C.c_void_p, C.c_double]
        L.orc_neighbor_pairs.restype = C.c_int64
        L.orc_neighbor_pairs.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.c_double, C.c_int, C.c_int64,
                                         C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_double),
                                         C.POINTER(C.c_int8)]
        L.orc_ik_compute.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.POINTER(OrcCG), C.c_double, C.c_int64,
                                     C.POINTER(C.c_double), C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.orc_ik_unique_bonds.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.POINTER(OrcCG), C.c_double, C.c_int64,
                                          C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.orc_thermo_compute.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.c_double, C.c_int, C.POINTER(C.c_double),
                                         C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                         C.POINTER(C.c_int64)]
        L.orc_gl_table.restype = C.c_int
        L.orc_gl_table.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.orc_num_threads.restype = C.c_int
        _lib = L
    return _lib


def ref_lib():
    """The reference's own potential/ and cgfunc/ headers (oracle/_ref). None when not built."""
    global _ref
    if _ref is None:
        build()
        path = os.path.join(_HERE, "_ref", "libpgt_refmath.so")
        if not os.path.exists(path):
            return None
        _ref = C.CDLL(path)
    return _ref


_refloops = None


def refloops_lib():
    """The reference's OWN calculators (IrvingKirkwood::compute, ThermoCalculator::compute, observables, CG functions,
    ParticleSystem, SimBox) compiled from /root/reference against the stand-ins of oracle/stubs/
    (oracle/ref_loops_shim.cpp -> oracle/_ref/libpgt_refloops.so).  None when not built."""
    global _refloops
    if _refloops is None:
        build()
        path = os.path.join(_HERE, "_ref", "libpgt_refloops.so")
        if not os.path.exists(path):
            return None
        lib()                                            # liborc.so first: the stand-in GSL rules resolve into it
        L = C.CDLL(path)
        L.ref_ik_compute.restype = C.c_int
        L.ref_thermo_compute.restype = C.c_int
        L.ref_cg_bond.restype = C.c_double
        _refloops = L
    return _refloops


def ref_ik(box, pos, diam, P, cg_kernel, cg_rcut, order, grid, vel=None, mass=None, obs_mask=1, dim=2, disp=None):
    """IrvingKirkwood::compute (irvingkirkwood/IrvingKirkwood.h:74-110) of the reference itself on `grid` (G x 3); dim = 3 uses
    the reference's 3-D typedefs (G x 9 fields, box periodic on all axes); obs_mask bit3 with `disp`: the elastic virial field."""
    L = refloops_lib()
    box = np.ascontiguousarray(np.asarray(box, np.float64)[:3])
    pos = np.ascontiguousarray(pos, np.float64)
    diam = np.ascontiguousarray(diam, np.float64)
    vel = None if vel is None else np.ascontiguousarray(vel, np.float64)
    mass = None if mass is None else np.ascontiguousarray(mass, np.float64)
    grid = np.ascontiguousarray(grid, np.float64)
    prm = np.ascontiguousarray(pot_params(P), np.float64)
    G = len(grid)
    D2 = dim * dim
    disp = None if disp is None else np.ascontiguousarray(disp, np.float64)
    vir = np.zeros((G, D2)) if obs_mask & 1 else None
    kin = np.zeros((G, D2)) if obs_mask & 2 else None
    ela = np.zeros((G, D2)) if (obs_mask & 8) and disp is not None else None
    mr, nl = C.c_double(0.0), C.c_int64(0)
    vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    kid = {"octic": 0, "rect": 1, "mollifier": 2}[cg_kernel]
    fn = L.ref_ik_compute if dim == 2 else L.ref_ik_compute3
    rc = fn(vp(box), C.c_int64(len(diam)), vp(pos), vp(vel), vp(diam), vp(mass), vp(disp), C.c_int(P.model),
            C.c_double(P.scaled_rcut), vp(prm), C.c_int(len(prm)), C.c_int(kid), C.c_double(cg_rcut), C.c_int(order),
            C.c_int(0), C.c_double(0.0), C.c_double(0.0), C.c_int(0), C.c_int64(G), vp(grid), C.c_uint(obs_mask),
            vp(vir), vp(kin), vp(ela), C.byref(mr), C.byref(nl))
    if rc != 0:
        raise RuntimeError("ref_ik_compute failed (%d)" % rc)
    return dict(virial=vir, kinetic=kin, elastic=ela, max_rcut=mr.value, n_live=nl.value)


def ref_thermo(box, pos, diam, P, vel=None, mass=None, obs_mask=1 | 4 | 8, dim=2, disp=None):
    """ThermoCalculator::compute (thermo/ThermoCalculator.h:68-90) of the reference itself; obs_mask bit0 virial,
    bit1 kinetic, bit2 energy, bit3 Born tensor."""
    L = refloops_lib()
    box = np.ascontiguousarray(np.asarray(box, np.float64)[:3])
    pos = np.ascontiguousarray(pos, np.float64)
    diam = np.ascontiguousarray(diam, np.float64)
    vel = None if vel is None else np.ascontiguousarray(vel, np.float64)
    mass = None if mass is None else np.ascontiguousarray(mass, np.float64)
    prm = np.ascontiguousarray(pot_params(P), np.float64)
    D2, D4 = dim * dim, dim ** 4
    out = np.zeros(2 * D2 + 1 + D4 + D2)
    disp = None if disp is None else np.ascontiguousarray(disp, np.float64)
    mr = C.c_double(0.0)
    vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    fn = L.ref_thermo_compute if dim == 2 else L.ref_thermo_compute3
    rc = fn(vp(box), C.c_int64(len(diam)), vp(pos), vp(vel), vp(diam), vp(mass), vp(disp), C.c_int(P.model),
            C.c_double(P.scaled_rcut), vp(prm), C.c_int(len(prm)), C.c_uint(obs_mask), vp(out), C.byref(mr))
    if rc != 0:
        raise RuntimeError("ref_thermo_compute failed (%d)" % rc)
    return dict(virial=out[0:D2].copy(), kinetic=out[D2:2 * D2].copy(), energy=float(out[2 * D2]),
                born=out[2 * D2 + 1:2 * D2 + 1 + D4].copy(), elastic=out[2 * D2 + 1 + D4:].copy(), max_rcut=mr.value)


# ---------------------------------------------------------------------------
# potential factories: coefficients exactly as pyglasstools/potential/__init__.py:78-143
def potential(name, **kw):
    P = OrcPot()
    P.model = MODEL_IDS[name]
    if name == "poly12":
        v0, eps, rc = kw.get("v0", 1.0), kw.get("eps", 0.2), kw.get("rcut", 1.25)
        p = [v0, eps, -28.0 * v0 / rc**12, 48.0 * v0 / rc**14, -21.0 * v0 / rc**16]
    elif name == "poly18":
        v0, eps, rc = kw.get("v0", 1.0), kw.get("eps", 0.0), kw.get("rcut", 1.25)
        p = [v0, eps, -55.0 * v0 / rc**18, 99.0 * v0 / rc**20, -45.0 * v0 / rc**22]
    elif name == "polylj":
        v0, eps, rc = kw.get("v0", 1.0), kw.get("eps", 0.2), kw.get("rcut", 2.5)
        p = [v0, eps, -28.0 * v0 / rc**12 + 10.0 / rc**6, 48.0 * v0 / rc**14 - 15.0 / rc**8,
             -21.0 * v0 / rc**16 + 6.0 / rc**10]
    elif name == "poly10":
        v0, eps, rc = kw.get("v0", 1.0), kw.get("eps", 0.0416667), kw.get("rcut", 1.48)
        p = [v0, eps, -(56.0) * v0 / (rc**10), (140.0) * v0 / (rc**12), -(120.0) * v0 / (rc**14), (35.0) * v0 / (rc**16)]
    elif name == "poly106":
        v0, eps, rc = kw.get("v0", 1.0), kw.get("eps", 0.1), kw.get("rcut", 3.0)
        p = [v0, eps, (-21 + 10 * rc**4) * v0 / rc**10, -((5 * (-7 + 3 * rc**4) * v0) / rc**12),
             (3 * (-5 + 2 * rc**4) * v0) / rc**14]
    elif name == "polyyukawa":
        v0, eps, rc, kappa = kw.get("v0", 10.0), kw.get("eps", 0.0), kw.get("rcut", 4.0), kw.get("kappa", 1.0)
        e = np.exp(-kappa * rc)
        p = [v0, kappa, eps, -((e * (15 + 7 * kappa * rc + kappa**2 * rc**2) * v0) / (8 * rc)),
             (e * (5 + 5 * kappa * rc + kappa**2 * rc**2) * v0) / (4 * rc**3),
             -((e * (3 + 3 * kappa * rc + kappa**2 * rc**2) * v0) / (8 * rc**5))]
    elif name in ("lj", "fslj"):
        rc = kw.get("rcut", 2.5)
        p = [kw.get("eps", 1.0), kw.get("vareps", 0.0)]
    else:
        raise KeyError(name)
    P.scaled_rcut = float(rc)
    for k, v in enumerate(p):
        P.p[k] = float(v)
    P.nparams = len(p)
    return P


def pot_params(P):
    n = getattr(P, "nparams", 8)
    return [P.p[k] for k in range(n)]


def cgfunc(name, rcut, order):
    cg = OrcCG()
    lib().orc_cg_init(C.byref(cg), KERNEL_IDS[name], C.c_double(rcut), int(order))
    return cg


def cgfunc_adaptive(name, rcut, relerr=1e-9, abserr=1e-9, maxiter=100):
    """cgfunc.<kernel>(rcut, mode="adaptive", relerr=, abserr=, maxiter=) -> CGFunc*Adaptive(maxiter, relerr, abserr, rcut)
    (cgfunc/__init__.py:37-44): GSL QAG with the 15-point Gauss-Kronrod rule."""
    cg = OrcCG()
    L = lib()
    L.orc_cg_init_adaptive.argtypes = [C.POINTER(OrcCG), C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
    L.orc_cg_init_adaptive.restype = None
    L.orc_cg_init_adaptive(C.byref(cg), KERNEL_IDS[name], rcut, relerr, abserr, int(maxiter))
    return cg


def qag_test(which, a, b, epsabs, epsrel, limit):
    """The QAG (GK15) restatement on a closed-form integrand: returns (value, status, abserr)."""
    L = lib()
    L.orc_qag_test.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.POINTER(C.c_int),
                               C.POINTER(C.c_double)]
    L.orc_qag_test.restype = C.c_double
    st, err = C.c_int(0), C.c_double(0.0)
    v = L.orc_qag_test(which, a, b, epsabs, epsrel, limit, C.byref(st), C.byref(err))
    return v, st.value, err.value


def gk15_tables():
    L = lib()
    xgk, wg, wgk = np.empty(8), np.empty(4), np.empty(8)
    L.orc_gk15_tables.argtypes = [C.c_void_p] * 3
    L.orc_gk15_tables.restype = None
    L.orc_gk15_tables(xgk.ctypes.data, wg.ctypes.data, wgk.ctypes.data)
    return xgk, wg, wgk


def pair_table(P, rij, di, dj):
    rij = np.ascontiguousarray(rij, np.float64)
    di = np.ascontiguousarray(di, np.float64)
    dj = np.ascontiguousarray(dj, np.float64)
    n = len(di)
    out = [np.empty(n) for _ in range(4)]
    lib().orc_pair_table(C.byref(P), C.c_int64(n), _dp(rij), _dp(di), _dp(dj), *[_dp(o) for o in out])
    return dict(energy=out[0], fdivr=out[1], stiff=out[2], rcutsq=out[3])


def ref_pair_table(P, rij, di, dj):
    R = ref_lib()
    rij = np.ascontiguousarray(rij, np.float64)
    di = np.ascontiguousarray(di, np.float64)
    dj = np.ascontiguousarray(dj, np.float64)
    n = len(di)
    out = [np.empty(n) for _ in range(4)]
    prm = np.array(pot_params(P), np.float64)
    rc = R.ref_pair_table(P.model, C.c_double(P.scaled_rcut), _dp(prm), len(prm), C.c_int64(n), _dp(rij), _dp(di),
                          _dp(dj), *[_dp(o) for o in out])
    assert rc == 0
    return dict(energy=out[0], fdivr=out[1], stiff=out[2], rcutsq=out[3])


def cg_table(cg, d2):
    d2 = np.ascontiguousarray(d2, np.float64)
    return np.array([lib().orc_cg_compute(C.byref(cg), C.c_double(v)) for v in d2])


def ref_cg_table(kernel, rcut, d2):
    d2 = np.ascontiguousarray(d2, np.float64)
    out = np.empty(len(d2))
    rc = ref_lib().ref_cg_table(KERNEL_IDS[kernel], C.c_double(rcut), C.c_int64(len(d2)), _dp(d2), _dp(out))
    assert rc == 0
    return out


def delta(cg, x, ri):
    x = np.ascontiguousarray(x, np.float64)
    ri = np.ascontiguousarray(ri, np.float64)
    return lib().orc_delta(C.byref(cg), _dp(x), _dp(ri))


def bond(cg, x, ri, rij):
    x, ri, rij = (np.ascontiguousarray(a, np.float64) for a in (x, ri, rij))
    return lib().orc_bond(C.byref(cg), _dp(x), _dp(ri), _dp(rij))


def gl_table(n):
    m = (n + 1) // 2
    x, w = np.empty(m), np.empty(m)
    lib().orc_gl_table(n, _dp(x), _dp(w))
    return x, w


class System:
    """ParticleSystem + SimBox as the reference builds them (utils.py:41-69)."""

    def __init__(self, box, pos, diam, vel=None, mass=None, dim=2, cell_hint=1.0):
        box = np.ascontiguousarray(np.asarray(box, np.float64)[:3])
        pos = np.ascontiguousarray(pos, np.float64)      # float32 inputs promote element-wise like pybind
        diam = np.ascontiguousarray(diam, np.float64)
        vel = None if vel is None else np.ascontiguousarray(vel, np.float64)
        mass = None if mass is None else np.ascontiguousarray(mass, np.float64)
        self.h = lib().orc_system_create(dim, _dp(box), C.c_int64(len(diam)), _dp(pos), _dp(vel), _dp(diam),
                                         _dp(mass), C.c_double(cell_hint))
        self.dim = dim
        self.n = lib().orc_system_n(self.h)
        self.vol = lib().orc_system_vol(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_system_destroy(self.h)
            self.h = None

    def max_rcut(self, P):
        return lib().orc_max_rcut(self.h, C.c_double(P.scaled_rcut))

    def wrapped(self):
        pos = np.empty((self.n, 3))
        ids = np.empty(self.n, np.int64)
        lib().orc_system_get(self.h, _dp(pos), _ip(ids))
        return pos, ids

    def neighbor_pairs(self, P, max_rcut=None, brute=False):
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        n = lib().orc_neighbor_pairs(self.h, C.byref(P), C.c_double(mr), int(brute), C.c_int64(0), None, None, None, None)
        pi, pj = np.empty(n, np.int64), np.empty(n, np.int64)
        rij = np.empty((n, 3))
        fl = np.empty(n, np.int8)
        lib().orc_neighbor_pairs(self.h, C.byref(P), C.c_double(mr), int(brute), C.c_int64(n), _ip(pi), _ip(pj), _dp(rij),
                                 fl.ctypes.data_as(C.POINTER(C.c_int8)))
        return pi, pj, rij, fl

    def ik(self, P, cg, grid, obs_mask=1, kinetic_mode=0, brute=False, max_rcut=None, counts=False):
        grid = np.ascontiguousarray(grid, np.float64)
        G = len(grid)
        D2 = self.dim * self.dim
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        vir = np.zeros((G, D2)) if obs_mask & 1 else None
        kin = np.zeros((G, D2)) if obs_mask & 2 else None
        rho = np.zeros(G) if obs_mask & 4 else None
        cnt = np.zeros((G, 3), np.int64) if counts else None
        lib().orc_ik_compute(self.h, C.byref(P), C.byref(cg), C.c_double(mr), C.c_int64(G), _dp(grid), obs_mask,
                             kinetic_mode, int(brute), _dp(vir), _dp(kin), _dp(rho), _ip(cnt))
        return dict(virial=vir, kinetic=kin, density=rho, counts=cnt)

    def ik_unique_bonds(self, P, cg, grid, max_rcut=None):
        grid = np.ascontiguousarray(grid, np.float64)
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        U = np.zeros(len(grid), np.int64)
        lib().orc_ik_unique_bonds(self.h, C.byref(P), C.byref(cg), C.c_double(mr), C.c_int64(len(grid)), _dp(grid), _ip(U))
        return U

    def thermo(self, P, brute=False, max_rcut=None, born=True):
        D = self.dim
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        vir, kin, bt, en = np.zeros(D * D), np.zeros(D * D), np.zeros(D**4), np.zeros(1)
        npairs = np.zeros(1, np.int64)
        lib().orc_thermo_compute(self.h, C.byref(P), C.c_double(mr), int(brute), _dp(vir), _dp(kin),
                                 _dp(bt) if born else None, _dp(en), _ip(npairs))
        return dict(virial=vir, kinetic=kin, born=bt, energy=en[0], npairs=int(npairs[0]))


    def set_displacement(self, disp):
        """ParticleSystem::setDisplacement: N x 3 rows in ORIGINAL particle order (None clears)."""
        L = lib()
        L.orc_system_set_displacement.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_system_set_displacement.restype = None
        d = None if disp is None else np.ascontiguousarray(disp, np.float64)
        L.orc_system_set_displacement(self.h, None if d is None else d.ctypes.data)

    def thermo_elastic(self, P, max_rcut=None):
        """GlobalElasticVirialStress (ListofCommonObservables.h:86-123 in ThermoCalculator::compute), times 1/vol."""
        L = lib()
        L.orc_thermo_elastic.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.c_double, C.c_void_p]
        L.orc_thermo_elastic.restype = None
        out = np.zeros(self.dim * self.dim)
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        L.orc_thermo_elastic(self.h, C.byref(P), mr, out.ctypes.data)
        return out

    def ik_elastic(self, P, cg, grid, max_rcut=None):
        """ElasticVirialStressField through IrvingKirkwood::compute."""
        L = lib()
        L.orc_ik_elastic.argtypes = [C.c_void_p, C.POINTER(OrcPot), C.POINTER(OrcCG), C.c_double, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_ik_elastic.restype = None
        grid = np.ascontiguousarray(grid, np.float64)
        out = np.zeros((len(grid), self.dim * self.dim))
        mr = self.max_rcut(P) if max_rcut is None else max_rcut
        L.orc_ik_elastic(self.h, C.byref(P), C.byref(cg), mr, len(grid), grid.ctypes.data, out.ctypes.data)
        return out


def cartesian_grid(L, dx):
    """ikcalculator grid (irvingkirkwood/__init__.py:45-53): i-major, both ends included."""
    nmax = int(L / dx)
    pts = np.linspace(-L / 2.0, +L / 2.0, nmax)
    g = np.zeros((nmax * nmax, 3))
    g[:, 0] = np.repeat(pts, nmax)
    g[:, 1] = np.tile(pts, nmax)
    return g, nmax


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))
