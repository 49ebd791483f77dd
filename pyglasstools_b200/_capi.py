"""ctypes binding of libpgt_cuda.so (include/pgt_cuda.h).

This is the only door from Python into the CUDA library.  There is no fallback: if the shared
library is missing or no CUDA device is present, importing / creating a context raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpgt_cuda.so")

POT_IDS = {"lj": 0, "forceshifted-lj": 1, "poly12": 2, "poly18": 3, "poly10": 4, "poly106": 5, "polyyukawa": 6,
           "polylj": 7}
CG_IDS = {"octic": 0, "rect": 1, "mollifier": 2}
OBS_VIRIAL, OBS_KINETIC, OBS_DENSITY, OBS_ENERGY, OBS_BORN, OBS_ELASTIC = 1, 2, 4, 8, 16, 32
FIELD_VIRIAL, FIELD_KINETIC, FIELD_DENSITY, FIELD_ELASTIC = 0, 1, 2, 3
STAGES = {"upload": 0, "cells": 1, "bonds": 2, "ik": 3, "thermo": 4, "download": 5}

EXPORTED = [
    "pgt_ctx_create", "pgt_ctx_destroy", "pgt_last_error", "pgt_device_count", "pgt_synchronize", "pgt_stream",
    "pgt_ik_download_sym_async", "pgt_frame_set_displacement", "pgt_thermo_elastic", "pgt_potential_set", "pgt_cg_set", "pgt_cg_set_adaptive", "pgt_cg_bond_adaptive", "pgt_set_max_rcut", "pgt_set_cell_size", "pgt_frame_upload",
    "pgt_frame_set_device", "pgt_build_cells", "pgt_build_bonds", "pgt_frame_counts", "pgt_grid_set_cartesian",
    "pgt_grid_set_points", "pgt_grid_set_shard", "pgt_ik_compute", "pgt_ik_download", "pgt_ik_download_async", "pgt_ik_download_wait", "pgt_ik_device_ptr",
    "pgt_thermo_compute", "pgt_export_bonds", "pgt_export_neighbors", "pgt_export_wrapped", "pgt_stage_time_ms",
    "pgt_kernel_launches", "pgt_pair_eval", "pgt_cg_delta", "pgt_cg_bond", "pgt_fp64_peak_probe",
    "pgt_ik_count_work", "pgt_event_record", "pgt_event_elapsed_ms", "pgt_host_alloc", "pgt_host_free",
    "pgt_set_option", "pgt_comm_unique_id", "pgt_comm_init", "pgt_comm_allgather_field", "pgt_comm_allreduce_sum", "pgt_comm_destroy",
    "pgt_frame_upload_part", "pgt_comm_allgather_frame", "pgt_thermo_set_shard", "pgt_ik_download_wait_keep", "pgt_thermo_compute_static",
    "pgt_frame_upload_xy", "pgt_thermo_compute_async", "pgt_host_register", "pgt_host_unregister",
]

_lib = None


def device_count():
    """CUDA devices visible to this process (0 without a driver)."""
    n = C.c_int(0)
    lib().pgt_device_count(C.byref(n))
    return n.value


def cg_bond_adaptive(kernel, rcut, relerr, abserr, maxiter, x, ri, rij):
    """AdaptiveCGFunc::getBondFunc on the host side of the library: (value, QAG exit code)."""
    kid = CG_IDS[kernel] if isinstance(kernel, str) else int(kernel)
    a = [np.ascontiguousarray(v, np.float64) for v in (x, ri, rij)]
    st = C.c_int(0)
    v = lib().pgt_cg_bond_adaptive(kid, float(rcut), float(abserr), float(relerr), int(maxiter), a[0].ctypes.data,
                                   a[1].ctypes.data, a[2].ctypes.data, C.byref(st))
    return v, st.value


class PgtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libpgt_cuda error %d: %s" % (code, msg))
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not found: build it with `make -C pyglasstools_b200/csrc` "
                              "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.pgt_last_error.restype = C.c_char_p
        L.pgt_last_error.argtypes = [C.c_void_p]
        L.pgt_stream.restype = C.c_void_p
        L.pgt_stream.argtypes = [C.c_void_p]
        L.pgt_kernel_launches.restype = C.c_int64
        L.pgt_kernel_launches.argtypes = [C.c_void_p]
        L.pgt_pair_eval.restype = C.c_double
        L.pgt_pair_eval.argtypes = [C.c_int, C.c_double, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double),
                                    C.c_double, C.c_double, C.c_int]
        L.pgt_cg_delta.restype = C.c_double
        L.pgt_cg_delta.argtypes = [C.c_int, C.c_double, C.c_double]
        L.pgt_cg_bond.restype = C.c_double
        L.pgt_cg_bond.argtypes = [C.c_int, C.c_double, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                  C.POINTER(C.c_double)]
        L.pgt_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.pgt_ctx_destroy.argtypes = [C.c_void_p]
        L.pgt_synchronize.argtypes = [C.c_void_p]
        L.pgt_potential_set.argtypes = [C.c_void_p, C.c_int, C.c_double, C.POINTER(C.c_double), C.c_int]
        L.pgt_cg_set.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_int]
        L.pgt_cg_set_adaptive.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
        L.pgt_cg_bond_adaptive.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.POINTER(C.c_int)]
        L.pgt_cg_bond_adaptive.restype = C.c_double
        L.pgt_set_max_rcut.argtypes = [C.c_void_p, C.c_double]
        L.pgt_set_cell_size.argtypes = [C.c_void_p, C.c_double]
        L.pgt_frame_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64,
                                       C.POINTER(C.c_double), C.c_int]
        L.pgt_frame_set_device.argtypes = L.pgt_frame_upload.argtypes
        L.pgt_frame_upload_xy.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64,
                                          C.POINTER(C.c_double)]
        L.pgt_build_cells.argtypes = [C.c_void_p]
        L.pgt_build_bonds.argtypes = [C.c_void_p]
        L.pgt_frame_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.pgt_grid_set_cartesian.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.c_int]
        L.pgt_grid_set_points.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int64]
        L.pgt_grid_set_shard.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
        L.pgt_ik_compute.argtypes = [C.c_void_p, C.c_uint]
        L.pgt_ik_download.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.pgt_ik_device_ptr.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.pgt_thermo_compute.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_double)]
        L.pgt_thermo_compute_static.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_double)]
        L.pgt_thermo_compute_async.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        L.pgt_thermo_elastic.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.pgt_frame_set_displacement.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int64]
        L.pgt_export_bonds.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p]
        L.pgt_export_neighbors.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
        L.pgt_export_wrapped.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.pgt_stage_time_ms.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float)]
        L.pgt_fp64_peak_probe.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.pgt_ik_count_work.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.pgt_event_record.argtypes = [C.c_void_p, C.c_int]
        L.pgt_event_elapsed_ms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]
        L.pgt_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
        L.pgt_host_free.argtypes = [C.c_void_p]
        L.pgt_host_register.argtypes = [C.c_void_p, C.c_uint64]
        L.pgt_host_unregister.argtypes = [C.c_void_p]
        L.pgt_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
        L.pgt_comm_unique_id.argtypes = [C.c_void_p]
        L.pgt_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.pgt_comm_allgather_field.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.c_void_p]
        L.pgt_comm_allreduce_sum.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int]
        L.pgt_comm_destroy.argtypes = [C.c_void_p]
        L.pgt_frame_upload_part.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64,
                                            C.POINTER(C.c_double), C.c_int]
        L.pgt_comm_allgather_frame.argtypes = [C.c_void_p]
        L.pgt_thermo_set_shard.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.pgt_ik_download_wait_keep.argtypes = [C.c_void_p, C.c_int]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Context:
    """One device, one stream (pgt_ctx)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().pgt_ctx_create(int(device), C.byref(self._h))
        if rc != 0:
            raise PgtError(rc, lib().pgt_last_error(None).decode())
        self.device = device
        self._keep = {}

    def close(self):
        if getattr(self, "_h", None) and self._h:
            lib().pgt_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise PgtError(rc, lib().pgt_last_error(self._h).decode())

    # models -----------------------------------------------------------------
    def set_potential(self, model, scaled_rcut, params):
        mid = POT_IDS[model] if isinstance(model, str) else int(model)
        p = np.ascontiguousarray(params, np.float64)
        self._ck(lib().pgt_potential_set(self._h, mid, float(scaled_rcut), _dp(p), len(p)))

    def set_cg(self, kernel, rcut, order, mode=0):
        kid = CG_IDS[kernel] if isinstance(kernel, str) else int(kernel)
        self._ck(lib().pgt_cg_set(self._h, kid, float(rcut), int(mode), int(order)))

    def set_cg_adaptive(self, kernel, rcut, relerr=1e-9, abserr=1e-9, maxiter=100):
        """cgfunc.<kernel>(rcut, mode="adaptive", ...): GSL QAG / 15-point Gauss-Kronrod bond quadrature."""
        kid = CG_IDS[kernel] if isinstance(kernel, str) else int(kernel)
        self._ck(lib().pgt_cg_set_adaptive(self._h, kid, float(rcut), float(abserr), float(relerr), int(maxiter)))

    def set_max_rcut(self, v):
        self._ck(lib().pgt_set_max_rcut(self._h, float(v)))

    def set_cell_size(self, v):
        self._ck(lib().pgt_set_cell_size(self._h, float(v)))

    # frame ------------------------------------------------------------------
    def upload(self, pos, diam, box, mass=None, vel=None, dim=2):
        f32 = pos.dtype == np.float32
        dt = np.float32 if f32 else np.float64
        pos = np.ascontiguousarray(pos, dt)
        diam = np.ascontiguousarray(diam, dt)
        mass = None if mass is None else np.ascontiguousarray(mass, dt)
        vel = None if vel is None else np.ascontiguousarray(vel, dt)
        self._keep["frame"] = (pos, diam, mass, vel)      # the copy is asynchronous
        b = (C.c_double * 3)(*[float(x) for x in box[:3]])
        vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        self._ck(lib().pgt_frame_upload(self._h, vp(pos), vp(diam), vp(mass), vp(vel), int(f32), len(diam), b, dim))
        self.dim = dim

    def upload_xy(self, pos_xy, diam, box, mass=None, vel=None):
        """pgt_frame_upload_xy: a 2-D frame from N x 2 position rows (z = 0)."""
        f32 = pos_xy.dtype == np.float32
        dt = np.float32 if f32 else np.float64
        pos_xy = np.ascontiguousarray(pos_xy, dt)
        assert pos_xy.ndim == 2 and pos_xy.shape[1] == 2
        diam = np.ascontiguousarray(diam, dt)
        mass = None if mass is None else np.ascontiguousarray(mass, dt)
        vel = None if vel is None else np.ascontiguousarray(vel, dt)
        self._keep["frame"] = (pos_xy, diam, mass, vel)   # the copy is asynchronous
        b = (C.c_double * 3)(*[float(x) for x in box[:3]])
        vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        self._ck(lib().pgt_frame_upload_xy(self._h, vp(pos_xy), vp(diam), vp(mass), vp(vel), int(f32), len(diam), b))
        self.dim = 2

    def upload_part(self, pos_part, diam_part, first, n_total, box, dim=2):
        """This rank's part of a frame (rows [first, first + len(diam_part)) of the N x 3 positions); assemble the frame
        on every GPU with comm_allgather_frame()."""
        f32 = pos_part.dtype == np.float32
        dt = np.float32 if f32 else np.float64
        pos_part = np.ascontiguousarray(pos_part, dt)
        diam_part = np.ascontiguousarray(diam_part, dt)
        self._keep["frame"] = (pos_part, diam_part)          # the copy is asynchronous
        b = (C.c_double * 3)(*[float(x) for x in box[:3]])
        self._ck(lib().pgt_frame_upload_part(self._h, pos_part.ctypes.data_as(C.c_void_p), diam_part.ctypes.data_as(C.c_void_p),
                                             int(f32), int(first), len(diam_part), int(n_total), b, dim))
        self.dim = dim

    def set_device_frame(self, d_pos, d_diam, n, box, is_f32=True, d_mass=None, d_vel=None, dim=2):
        b = (C.c_double * 3)(*[float(x) for x in box[:3]])
        self._ck(lib().pgt_frame_set_device(self._h, C.c_void_p(d_pos), C.c_void_p(d_diam),
                                            C.c_void_p(d_mass) if d_mass else None,
                                            C.c_void_p(d_vel) if d_vel else None, int(is_f32), int(n), b, dim))
        self.dim = dim

    def build_cells(self):
        self._ck(lib().pgt_build_cells(self._h))

    def build_bonds(self):
        self._ck(lib().pgt_build_bonds(self._h))

    def counts(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(lib().pgt_frame_counts(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(n_live=a.value, n_halfbonds=b.value, n_cells=c.value)

    # grid -------------------------------------------------------------------
    def set_grid_cartesian(self, xs, ys):
        xs = np.ascontiguousarray(xs, np.float64)
        ys = np.ascontiguousarray(ys, np.float64)
        self._ck(lib().pgt_grid_set_cartesian(self._h, _dp(xs), len(xs), _dp(ys), len(ys)))
        self.G = len(xs) * len(ys)
        self.g0, self.g1 = 0, self.G

    def set_grid_points(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
        self._ck(lib().pgt_grid_set_points(self._h, _dp(xyz), len(xyz)))
        self.G = len(xyz)
        self.g0, self.g1 = 0, self.G

    def set_shard(self, g0, g1):
        self._ck(lib().pgt_grid_set_shard(self._h, int(g0), int(g1)))
        self.g0, self.g1 = int(g0), int(g1)

    # compute ----------------------------------------------------------------
    def ik_compute(self, obs_mask=OBS_VIRIAL):
        self._ck(lib().pgt_ik_compute(self._h, int(obs_mask)))

    def ik_download(self, field=FIELD_VIRIAL, out=None):
        n = self.g1 - self.g0
        shape = (n,) if field == FIELD_DENSITY else (n, getattr(self, "dim", 2) ** 2)
        if out is None:
            out = np.empty(shape, np.float64)
        self._ck(lib().pgt_ik_download(self._h, int(field), out.ctypes.data_as(C.c_void_p)))
        return out

    def ik_download_async(self, field, out):
        """Enqueues the read-back into the page-locked array `out`; valid after sync()."""
        self._ck(lib().pgt_ik_download_async(self._h, int(field), out.ctypes.data_as(C.c_void_p)))
        return out

    def ik_download_sym_async(self, field, out):
        """Same for a 2x2 tensor field as (n, 3) rows (xx, xy, yy) -- the tensors are symmetric, the copy 25 % smaller."""
        self._ck(lib().pgt_ik_download_sym_async(self._h, int(field), out.ctypes.data_as(C.c_void_p)))
        return out

    def ik_download_wait(self):
        self._ck(lib().pgt_ik_download_wait(self._h))

    def ik_download_wait_keep(self, keep):
        """Waits until at most `keep` asynchronous read-backs are still in flight."""
        self._ck(lib().pgt_ik_download_wait_keep(self._h, int(keep)))

    def ik_device_ptr(self, field=FIELD_VIRIAL):
        p, n = C.c_void_p(), C.c_int64()
        self._ck(lib().pgt_ik_device_ptr(self._h, int(field), C.byref(p), C.byref(n)))
        return p.value, n.value

    def thermo(self, obs_mask=OBS_VIRIAL | OBS_ENERGY):
        D = getattr(self, "dim", 2)
        D2, D4 = D * D, D ** 4
        out = np.zeros(2 * D2 + 1 + D4, np.float64)
        self._ck(lib().pgt_thermo_compute(self._h, int(obs_mask), _dp(out)))
        return dict(virial=out[0:D2].copy(), kinetic=out[D2:2 * D2].copy(), energy=float(out[2 * D2]), born=out[2 * D2 + 1:].copy())

    def thermo_async(self, obs_mask, out):
        """pgt_thermo_compute_async: the 25 doubles (virial 4 | kinetic 4 | energy | born 16) land in `out` (a page-locked
        float64 array of at least 25 entries) once the context has been synchronised."""
        self._ck(lib().pgt_thermo_compute_async(self._h, int(obs_mask), out.ctypes.data_as(C.c_void_p)))

    def thermo_static(self, obs_mask=OBS_VIRIAL | OBS_ENERGY):
        """pgt_thermo_compute_static: the same sums with the library's static device scratch, no allocation in the call."""
        out = np.zeros(25, np.float64)
        self._ck(lib().pgt_thermo_compute_static(self._h, int(obs_mask), _dp(out)))
        return dict(virial=out[0:4].copy(), kinetic=out[4:8].copy(), energy=float(out[8]), born=out[9:25].copy())

    def set_displacement(self, disp):
        """ParticleSystem::setDisplacement: N x 3 rows in the uploaded frame's particle order."""
        d = np.ascontiguousarray(disp, np.float64).reshape(-1, 3)
        self._ck(lib().pgt_frame_set_displacement(self._h, _dp(d), C.c_int64(len(d))))

    def thermo_elastic(self):
        out = np.zeros(getattr(self, "dim", 2) ** 2, np.float64)
        self._ck(lib().pgt_thermo_elastic(self._h, _dp(out)))
        return out

    # exports ----------------------------------------------------------------
    def export_bonds(self):
        n = C.c_int64()
        self._ck(lib().pgt_export_bonds(self._h, 0, C.byref(n), None, None, None, None))
        m = n.value
        i, j = np.empty(m, np.int32), np.empty(m, np.int32)
        rij, f = np.empty((m, 2)), np.empty(m)
        if m:
            vp = lambda a: a.ctypes.data_as(C.c_void_p)
            self._ck(lib().pgt_export_bonds(self._h, m, C.byref(n), vp(i), vp(j), vp(rij), vp(f)))
        return i, j, rij, f

    def export_neighbors(self):
        n = C.c_int64()
        self._ck(lib().pgt_export_neighbors(self._h, 0, C.byref(n), None, None))
        m = n.value
        i, j = np.empty(m, np.int32), np.empty(m, np.int32)
        if m:
            vp = lambda a: a.ctypes.data_as(C.c_void_p)
            self._ck(lib().pgt_export_neighbors(self._h, m, C.byref(n), vp(i), vp(j)))
        return i, j

    def export_wrapped(self, n):
        xy = np.empty((n, 2))
        alive = np.empty(n, np.uint8)
        self._ck(lib().pgt_export_wrapped(self._h, xy.ctypes.data_as(C.c_void_p), alive.ctypes.data_as(C.c_void_p)))
        return xy, alive.astype(bool)

    # instrumentation ----------------------------------------------------------
    def stage_ms(self, stage):
        ms = C.c_float()
        self._ck(lib().pgt_stage_time_ms(self._h, STAGES[stage] if isinstance(stage, str) else stage, C.byref(ms)))
        return ms.value

    def sync(self):
        self._ck(lib().pgt_synchronize(self._h))

    def stream(self):
        return lib().pgt_stream(self._h)

    def launches(self):
        return lib().pgt_kernel_launches(self._h)

    def ik_count_work(self):
        u, it, pa = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(lib().pgt_ik_count_work(self._h, C.byref(u), C.byref(it), C.byref(pa)))
        return dict(unique_bonds=u.value, halfbond_items=it.value, particles=pa.value)

    def event_record(self, slot):
        self._ck(lib().pgt_event_record(self._h, int(slot)))

    def event_elapsed_ms(self, a, b):
        ms = C.c_float()
        self._ck(lib().pgt_event_elapsed_ms(self._h, int(a), int(b), C.byref(ms)))
        return ms.value

    def set_option(self, name, value):
        self._ck(lib().pgt_set_option(self._h, name.encode(), float(value)))

    def fp64_peak(self):
        t = C.c_double()
        self._ck(lib().pgt_fp64_peak_probe(self._h, C.byref(t)))
        return t.value

    # nccl -------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * 128)()
        rc = lib().pgt_comm_unique_id(buf)
        if rc != 0:
            raise PgtError(rc, lib().pgt_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, nranks, rank, uid):
        buf = (C.c_char * 128).from_buffer_copy(uid)
        self._ck(lib().pgt_comm_init(self._h, nranks, rank, buf))

    def comm_allgather_field(self, field, counts, out_dev_ptr):
        c = (C.c_int64 * len(counts))(*[int(x) for x in counts])
        self._ck(lib().pgt_comm_allgather_field(self._h, int(field), c, C.c_void_p(out_dev_ptr)))

    def comm_allgather_frame(self):
        self._ck(lib().pgt_comm_allgather_frame(self._h))

    def thermo_set_shard(self, part, nparts):
        self._ck(lib().pgt_thermo_set_shard(self._h, int(part), int(nparts)))

    def comm_allreduce_sum(self, arr):
        a = np.ascontiguousarray(arr, np.float64)
        self._ck(lib().pgt_comm_allreduce_sum(self._h, _dp(a), a.size))
        return a


def host_register(arr):
    """Page-locks a numpy array's memory in place (pgt_host_register); False when the driver refuses."""
    return lib().pgt_host_register(arr.ctypes.data_as(C.c_void_p), C.c_uint64(arr.nbytes)) == 0


def host_unregister(arr):
    return lib().pgt_host_unregister(arr.ctypes.data_as(C.c_void_p)) == 0


class PinnedArray:
    """A numpy view over page-locked host memory (cudaHostAlloc) owned by this object."""

    def __init__(self, shape, dtype):
        self.shape = tuple(int(x) for x in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = C.c_void_p()
        rc = lib().pgt_host_alloc(C.byref(self._p), max(nbytes, 1))
        if rc != 0:
            raise PgtError(rc, lib().pgt_last_error(None).decode())
        buf = (C.c_char * max(nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._p:
            self.array = None
            lib().pgt_host_free(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def pair_eval(model, scaled_rcut, params, rij, di, dj, what):
    p = np.ascontiguousarray(params, np.float64)
    r = np.ascontiguousarray(rij, np.float64)
    mid = POT_IDS[model] if isinstance(model, str) else int(model)
    return lib().pgt_pair_eval(mid, float(scaled_rcut), _dp(p), len(p), _dp(r), float(di), float(dj), int(what))


def cg_delta(kernel, rcut, dr_sq):
    return lib().pgt_cg_delta(CG_IDS[kernel] if isinstance(kernel, str) else int(kernel), float(rcut), float(dr_sq))


def cg_bond(kernel, rcut, order, x, ri, rij):
    x, ri, rij = (np.ascontiguousarray(a, np.float64) for a in (x, ri, rij))
    return lib().pgt_cg_bond(CG_IDS[kernel] if isinstance(kernel, str) else int(kernel), float(rcut), int(order),
                             _dp(x), _dp(ri), _dp(rij))
