"""Pair potentials (reference: pyglasstools/potential/__init__.py:9-143).

Each factory computes the smoothing coefficients c0..c3 that make the potential and its first two (three)
derivatives vanish at the cut-off and hands ``(scaled_rcut, params)`` to the C++ class; constructing a
potential registers it as the module-level current potential, as in the reference.
"""
import numpy as np

import pyglasstools
from pyglasstools.potential import _potential


class pairpotential(object):
    """Common accessors of every pair potential."""

    def __init__(self):
        self.cpppairpotential = None
        self.name = None
        self.dict_params = None

    def _install(self, name, cppclass, rcut, params, labels):
        self.name = name
        self.cpppairpotential = cppclass(rcut, [float(p) for p in params])
        self.dict_params = {label: k for k, label in enumerate(labels)}
        pyglasstools.set_potential(self)

    def get_potentialname(self):
        return self.name

    def set_scaledrcut(self, rcut):
        self.cpppairpotential.scaled_rcut = rcut

    def get_scaledrcut(self):
        return self.cpppairpotential.scaled_rcut

    def _slot(self, num):
        return self.dict_params[num] if isinstance(num, str) else num

    def set_params(self, val, num):
        if isinstance(num, (str, int)):
            self.cpppairpotential.setParams(val, self._slot(num))

    def get_params(self, num=None):
        params = self.cpppairpotential.getParams()
        if num is None:
            return params
        if isinstance(num, (str, int)):
            return params[self._slot(num)]

    def print_params(self):
        params = self.cpppairpotential.getParams()
        for key, slot in self.dict_params.items():
            print("{}: {} ".format(key, params[slot]))

    @staticmethod
    def _vec3(rij):
        rij = np.asarray(rij, dtype=np.float64)
        return np.append(rij, 0.0) if len(rij) == 2 else rij

    def get_pairforce(self, rij, di, dj):
        r = self._vec3(rij)
        return self.cpppairpotential.getPairForceDivR(r, di, dj) * np.linalg.norm(r)

    def get_pairenergy(self, rij, di, dj):
        return self.cpppairpotential.getPairEnergy(self._vec3(rij), di, dj)


class lj(pairpotential):
    """Lennard-Jones, truncated-and-shifted or force-shifted."""

    def __init__(self, eps, vareps, rcut, name="lennard-jones", mode=None):
        classes = {"truncated": _potential.PairPotentialLJ, "force-shifted": _potential.PairPotentialForceShiftedLJ}
        if mode is None:
            raise NameError('Please select a Lennard-Jones potential available modes: truncated and force-shifted are available')
        if mode not in classes:
            raise NameError('Lennard jones potential mode not recognized. Only: truncated and force-shifted are available')
        self._install("{}+{}".format(name, mode), classes[mode], rcut, [eps, vareps], ['eps', 'vareps'])
        self.dict_params = {'eps': 0, 'vareps': 0}      # sic: the reference maps both labels to slot 0


_POLY_LABELS = ['v0', 'eps', 'c0', 'c1', 'c2']


class polydisperse12(pairpotential):
    """v0 (sigma/r)^12 + c0 + c1 (r/sigma)^2 + c2 (r/sigma)^4"""

    def __init__(self, v0=1.0, eps=0.2, rcut=1.25):
        c = (-28.0 * v0 / rcut**12, 48.0 * v0 / rcut**14, -21.0 * v0 / rcut**16)
        self._install("polydisperse-12", _potential.PairPotentialPoly12, rcut, (v0, eps) + c, _POLY_LABELS)


class polydisperse18(pairpotential):
    """v0 (sigma/r)^18 + c0 + c1 (r/sigma)^2 + c2 (r/sigma)^4"""

    def __init__(self, v0=1.0, eps=0.0, rcut=1.25):
        c = (-55.0 * v0 / rcut**18, 99.0 * v0 / rcut**20, -45.0 * v0 / rcut**22)
        self._install("polydisperse-18", _potential.PairPotentialPoly18, rcut, (v0, eps) + c, _POLY_LABELS)


class polydisperselj(pairpotential):
    """v0 [(sigma/r)^12 - (sigma/r)^6] + smoothing polynomial"""

    def __init__(self, v0=1.0, eps=0.2, rcut=2.5):
        c = (-28.0 * v0 / rcut**12 + 10.0 / rcut**6, 48.0 * v0 / rcut**14 - 15.0 / rcut**8,
             -21.0 * v0 / rcut**16 + 6.0 / rcut**10)
        self._install("polydisperse-lj", _potential.PairPotentialPolyLJ, rcut, (v0, eps) + c, _POLY_LABELS)


class polydisperse10(pairpotential):
    """v0 (sigma/r)^10 + c0 + c1 (r/sigma)^2 + c2 (r/sigma)^4 + c3 (r/sigma)^6"""

    def __init__(self, v0=1.0, eps=0.0416667, rcut=1.48):
        c = (-(56.0) * v0 / (rcut**10), (140.0) * v0 / (rcut**12), -(120.0) * v0 / (rcut**14), (35.0) * v0 / (rcut**16))
        self._install("polydisperse-10", _potential.PairPotentialPoly10, rcut, (v0, eps) + c, _POLY_LABELS + ['c3'])


class polydisperse106(pairpotential):
    """v0 [(sigma/r)^10 - (sigma/r)^6] + smoothing polynomial"""

    def __init__(self, v0=1.0, eps=0.1, rcut=3.0):
        c = ((-21 + 10 * rcut**4) * v0 / rcut**10, -((5 * (-7 + 3 * rcut**4) * v0) / rcut**12),
             (3 * (-5 + 2 * rcut**4) * v0) / rcut**14)
        self._install("polydisperse-106", _potential.PairPotentialPoly106, rcut, (v0, eps) + c, _POLY_LABELS)


class polydisperseyukawa(pairpotential):
    """v0 exp(-kappa r/sigma) / (r/sigma) + smoothing polynomial"""

    def __init__(self, v0=10.0, eps=0.0, rcut=4.0, kappa=1.0):
        e = np.exp(-kappa * rcut)
        kr = kappa * rcut
        c = (-((e * (15 + 7 * kr + kappa**2 * rcut**2) * v0) / (8 * rcut)),
             (e * (5 + 5 * kr + kappa**2 * rcut**2) * v0) / (4 * rcut**3),
             -((e * (3 + 3 * kr + kappa**2 * rcut**2) * v0) / (8 * rcut**5)))
        self._install("polydisperse-yukawa", _potential.PairPotentialPolyYukawa, rcut, (v0, kappa, eps) + c,
                      ['v0', 'kappa', 'eps', 'c0', 'c1', 'c2'])
