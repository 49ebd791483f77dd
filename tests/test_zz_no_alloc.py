"""pgt_thermo_compute_static (include/pgt_cuda.h): the thermo pair sums with the library's file-scope device scratch, no
allocation in the call.  No counterpart in the reference; the checker is the
existing entry point, pgt_thermo_compute, on the same context and frame, through the C-ABI, at the tolerance of the other
GPU parity tests (BASELINE.json's 1e-10 relative / 1e-12 absolute per component).

The file is named so that pytest collects it last; it builds its own contexts and frames, so it passes alone in a fresh
process as well as after the other tests.
"""
import numpy as np
import pytest

from pyglasstools_b200 import _capi, synth

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-10, 1e-12
MAX_PARTICLES = 2097152          # PGT_THERMO_STATIC_MAX_PARTICLES
ALL = _capi.OBS_VIRIAL | _capi.OBS_ENERGY | _capi.OBS_BORN | _capi.OBS_KINETIC


def _vec(t):
    return np.concatenate([t["virial"], t["kinetic"], [t["energy"]], t["born"]])


def _close(got, want):
    got, want = _vec(got), _vec(want)
    err = np.abs(got - want)
    assert np.all(err <= ATOL + RTOL * np.abs(want)), "max abs err %.3e" % err.max()


def _load(ctx, n_side, seed, rc=1.25):
    fr = synth.synth_frame(n_side, seed, velocities=True)
    ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
    dmax = float(fr["diameter"].max())
    ctx.set_max_rcut(max(rc * dmax, dmax))
    ctx.upload(fr["position"], fr["diameter"], fr["box"], mass=fr.get("mass"), vel=fr.get("velocity"))
    ctx.build_cells()
    return fr


def test_static_scratch_thermo_equals_the_allocating_entry_point_twice_in_a_row():
    ctx = _capi.Context(0)
    try:
        # a frame that fills many blocks of partial sums
        _load(ctx, 160, 7)
        want = ctx.thermo(ALL)
        assert abs(want["energy"]) > 0 and np.abs(want["virial"]).max() > 0 and np.abs(want["born"]).max() > 0
        first = ctx.thermo_static(ALL)
        second = ctx.thermo_static(ALL)
        _close(first, want)
        _close(second, want)
        assert np.array_equal(_vec(first), _vec(second))
        # a narrower observable mask after the full one: nothing of the last call's result may show through
        v_want = ctx.thermo(_capi.OBS_VIRIAL)
        v_got = ctx.thermo_static(_capi.OBS_VIRIAL)
        _close(v_got, v_want)
        assert v_got["energy"] == 0.0 and not v_got["born"].any() and not v_got["kinetic"].any()
        # a smaller frame: fewer blocks than the scratch still holds partial sums for
        _load(ctx, 48, 11)
        want2 = ctx.thermo(ALL)
        _close(ctx.thermo_static(ALL), want2)
        _close(ctx.thermo_static(ALL), want2)
        # a second context on the same device shares the array (one call after the other: no overlap on the GPU)
        other = _capi.Context(0)
        try:
            _load(other, 96, 5)
            want3 = other.thermo(ALL)
            _close(other.thermo_static(ALL), want3)
            _close(ctx.thermo_static(ALL), want2)
            _close(other.thermo_static(ALL), want3)
        finally:
            other.close()
    finally:
        ctx.close()


def test_static_scratch_thermo_refuses_frames_beyond_its_stated_maximum():
    ctx = _capi.Context(0)
    try:
        n = MAX_PARTICLES + 1
        rng = np.random.default_rng(3)
        L = float(np.sqrt(n))
        pos = np.zeros((n, 3), np.float32)
        pos[:, :2] = rng.uniform(-L / 2, L / 2, (n, 2)).astype(np.float32)
        diam = np.ones(n, np.float32)
        rc = 1.25
        ctx.set_potential("poly12", rc, [1.0, 0.2, -28.0 / rc**12, 48.0 / rc**14, -21.0 / rc**16])
        ctx.set_max_rcut(rc)
        ctx.upload(pos, diam, np.array([L, L, 1.0, 0.0, 0.0, 0.0]))
        ctx.build_cells()
        with pytest.raises(_capi.PgtError) as e:
            ctx.thermo_static(_capi.OBS_VIRIAL)
        assert "at most %d" % MAX_PARTICLES in str(e.value)
        # the context is still good: the limit itself is served
        pos, diam = pos[:MAX_PARTICLES], diam[:MAX_PARTICLES]
        ctx.upload(pos, diam, np.array([L, L, 1.0, 0.0, 0.0, 0.0]))
        ctx.build_cells()
        _close(ctx.thermo_static(_capi.OBS_VIRIAL | _capi.OBS_ENERGY), ctx.thermo(_capi.OBS_VIRIAL | _capi.OBS_ENERGY))
    finally:
        ctx.close()
