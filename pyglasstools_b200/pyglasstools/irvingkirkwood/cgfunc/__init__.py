"""Coarse-graining functions (reference: pyglasstools/irvingkirkwood/cgfunc/__init__.py:8-82).

``octic``, ``rect`` and ``mollifier`` take ``rcut`` and ``mode``: ``"fixedpoint"`` with ``order=`` (Gauss-Legendre
nodes of the bond-line integral, the rule GSL's ``glfixed`` applies) or ``"adaptive"`` with ``relerr=, abserr=,
maxiter=`` (GSL QAG in the reference; constructible here, but the device path rejects it).
"""
from pyglasstools.irvingkirkwood.cgfunc import _cgfunc


class basecgfunc(object):
    _fixed = None
    _adaptive = None

    def __init__(self, rcut=2.5, mode=None, **kwargs):
        if mode == "fixedpoint":
            self.cgfunc = self._fixed(kwargs["order"], rcut)
        elif mode == "adaptive":
            self.cgfunc = self._adaptive(kwargs["maxiter"], kwargs["relerr"], kwargs["abserr"], rcut)
        # any other mode leaves the object without a kernel, as the reference does

    def _getCGFunc(self):
        return self.cgfunc

    def set_rcut(self, rcut):
        self.cgfunc.setRcut(rcut)

    def get_rcut(self):
        return self.cgfunc.getRcut()

    def get_deltafunc(self, x, ri):
        return self.cgfunc.getDeltaFunc(x, ri)

    def get_objfunc(self, s, x, ri, rij):
        return self.cgfunc.getObjFunc(s, x, ri, rij)

    def get_bondfunc(self, x, ri, rij):
        return self.cgfunc.getBondFunc(x, ri, rij)


class octic(basecgfunc):
    """15/(8 pi R^2) (1 - (d/R)^4)^2 for d < R."""
    _fixed = _cgfunc.CGFuncOcticFixedPoint
    _adaptive = _cgfunc.CGFuncOcticAdaptive


class rect(basecgfunc):
    """1/(pi R^2) for d < R."""
    _fixed = _cgfunc.CGFuncRectFixedPoint
    _adaptive = _cgfunc.CGFuncRectAdaptive


class mollifier(basecgfunc):
    """C exp(-1/(1 - (d/R)^2)) for d < R."""
    _fixed = _cgfunc.CGFuncMollifierFixedPoint
    _adaptive = _cgfunc.CGFuncMollifierAdaptive
