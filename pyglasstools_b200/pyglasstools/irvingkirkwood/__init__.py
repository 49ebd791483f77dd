"""Irving-Kirkwood coarse-grained fields (reference: pyglasstools/irvingkirkwood/__init__.py:8-166).

Grid points are split over the ranks exactly as the reference does (``np.array_split`` chunks in i-major order,
first ``G mod size`` chunks one longer); one rank = one process = one GPU.  No message is needed for that: every
rank builds the (cheap) grid itself and keeps its chunk.
"""
import numpy as np

import pyglasstools
from pyglasstools import _pyglasstools, comm, rank, size, solvers_list
from pyglasstools.irvingkirkwood import _irvingkirkwood


def initialize_field(names, dim, gridpoints):
    """One field holder per observable family in ``names``, each sized for this rank's grid points."""
    if dim not in (2, 3):
        return {}
    suffix = "Field{}D".format(dim)
    npoints = len(gridpoints)
    found = {}

    def make(key, stem, islocal):
        field = getattr(_irvingkirkwood, stem + suffix)(key, "2-TENSOR", islocal, npoints)
        field.addGridpoints(gridpoints)
        found[key] = field

    if any("virialstress" in s for s in names):
        if any("elastic" in s for s in names):
            make("elasticvirialstress", "ElasticVirialStress", False)
        if not all("elastic" in s for s in names):
            make("virialstress", "VirialStress", False)
    if any("kineticstress" in s for s in names):
        make("kineticstress", "KineticStress", True)
    return found


def shard_bounds(total, nranks):
    """[g0, g1) of every rank under ``np.array_split(range(total), nranks)``."""
    base, extra = divmod(int(total), int(nranks))
    sizes = [base + 1 if r < extra else base for r in range(nranks)]
    starts = np.concatenate([[0], np.cumsum(sizes)])
    return [(int(starts[r]), int(starts[r + 1])) for r in range(nranks)]


class _ikbase(object):
    def _finish(self, cgfunc):
        self.calculator = _irvingkirkwood.IrvingKirkwood(pyglasstools.get_sysdata().cppparticledata,
                                                         pyglasstools.get_potential().cpppairpotential,
                                                         cgfunc._getCGFunc(), comm)
        solvers_list.append(self)

    def add_observables(self, observables):
        for name in observables:
            self.calculator.addObservable(observables[name])

    def run(self):
        self.calculator.compute(self.gridpoints)

    def update(self, frame_num):
        pyglasstools.update_sysdata(frame_num)
        self.calculator.setSystemData(pyglasstools.get_sysdata().cppparticledata)


class ikcalculator(_ikbase):
    """Stress fields on the square lattice ``linspace(-L/2, +L/2, int(L/dx))^2`` (both ends included)."""

    def __init__(self, cgfunc, dx):
        self.cgfunc = cgfunc
        self.manager = _pyglasstools.Manager()
        self.coordmode = "cartesian"
        L = pyglasstools.get_sysdata().pysimbox.boxsize[0]
        nmax = int(L / dx)
        points = np.linspace(-L / 2.0, +L / 2.0, nmax)
        self._axis = np.ascontiguousarray(points, dtype=np.float64)
        self.gridsize = nmax * nmax
        self._g0, self._g1 = shard_bounds(self.gridsize, size)[rank]
        g = np.arange(self._g0, self._g1)
        chunk = np.zeros((len(g), 3), dtype=np.float64)
        if nmax > 0:
            chunk[:, 0] = points[g // nmax]
            chunk[:, 1] = points[g % nmax]
        self.gridpoints = chunk
        self._chunk_copy = chunk.copy()          # the lattice as generated: an in-place edit of gridpoints is noticed
        self._finish(cgfunc)

    def run(self):
        gp = self.gridpoints
        if (isinstance(gp, np.ndarray) and gp.dtype == np.float64 and gp.flags.c_contiguous and gp.shape == self._chunk_copy.shape
                and _pyglasstools.array_equal_fast(gp, self._chunk_copy)):
            # same points, but the device generates them from the axis instead of reading G x 3 doubles
            self.calculator.compute_cartesian(self._axis, self._axis, self._g0, self._g1)
        else:
            self.calculator.compute(self.gridpoints)


class radialcalculator(_ikbase):
    """Stress fields on rings around ``center``: radii linear (``dr``) or logarithmic (``dlnr``), points on a
    ring spaced ``dr`` apart; the centre itself is point 0."""

    def __init__(self, cgfunc, center=np.array([0, 0, 0]), spacingtype="normal", rmax=10, rmin=0, dr=0.1, dlnr=0.05):
        self.cgfunc = cgfunc
        self.manager = _pyglasstools.Manager()
        self.coordmode = "polar"
        self.center = np.asarray(center, dtype=np.float64)
        if spacingtype == "normal":
            radii = np.linspace(0, rmax, int(rmax / dr))[1:]
        elif spacingtype == "log":
            radii = np.exp(np.linspace(np.log(rmin), np.log(rmax), int((np.log(rmax) - np.log(rmin)) / dlnr)))
        else:
            raise ValueError("spacingtype must be 'normal' or 'log'")
        ring_id, ring_r, ring_theta, pts = [0.0], [0.0], [0.0], [self.center.copy()]
        for k, r in enumerate(radii):
            theta = np.linspace(0, 2 * np.pi, int(2 * np.pi / (dr / r)))
            ring_id.extend([k + 1.0] * len(theta))
            ring_r.extend([r] * len(theta))
            ring_theta.extend(theta)
            xyz = np.zeros((len(theta), 3))
            xyz[:, 0] = r * np.cos(theta)
            xyz[:, 1] = r * np.sin(theta)
            pts.extend(xyz + self.center)
        self.gridsize = len(pts)
        g0, g1 = shard_bounds(self.gridsize, size)[rank]
        self.gridid = np.asarray(ring_id, dtype=np.float64)[g0:g1]
        self.globalr = np.asarray(ring_r, dtype=np.float64)[g0:g1]
        self.globaltheta = np.asarray(ring_theta, dtype=np.float64)[g0:g1]
        self.gridpoints = np.ascontiguousarray(np.asarray(pts, dtype=np.float64)[g0:g1]).reshape(g1 - g0, 3)
        self._finish(cgfunc)

    def set_center(self, center):
        self.center = center
        self.gridpoints[0:] += center
