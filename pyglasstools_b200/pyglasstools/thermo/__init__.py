"""Global (thermodynamic) observables (reference: pyglasstools/thermo/__init__.py:8-56)."""
import pyglasstools
from pyglasstools import _pyglasstools, solvers_list
from pyglasstools.thermo import _thermo

# keyword in the requested names -> (key in the returned dict, class stem, tensor type, islocal)
_GLOBAL_KINDS = (
    ("potentialenergy", "potentialenergy", "GlobalPotentialEnergy", "SCALAR", False),
    ("kineticstress", "kineticstress", "GlobalKineticStress", "2-TENSOR", True),
    ("borntensor", "borntensor", "GlobalBornTensor", "4-TENSOR", False),
)


def initialize_global(names, dim):
    """Builds one holder per observable family mentioned in ``names`` (e.g. 'virialstress_01')."""
    if dim not in (2, 3):
        return {}
    suffix = "{}D".format(dim)
    found = {}

    def make(key, stem, kind, islocal):
        found[key] = getattr(_thermo, stem + suffix)(key, kind, islocal)

    if any("potentialenergy" in s for s in names):
        make(*_GLOBAL_KINDS[0][1:])
    if any("virialstress" in s for s in names):
        if any("elastic" in s for s in names):
            make("elasticvirialstress", "GlobalElasticVirialStress", "2-TENSOR", False)
        else:
            make("virialstress", "GlobalVirialStress", "2-TENSOR", False)
    for keyword, key, stem, kind, islocal in _GLOBAL_KINDS[1:]:
        if any(keyword in s for s in names):
            make(key, stem, kind, islocal)
    return found


class calculator(object):
    """Computes the registered global observables of the current frame on the GPU."""

    def __init__(self):
        self.manager = _pyglasstools.Manager()
        self.thermocalculator = _thermo.ThermoCalculator(pyglasstools.get_sysdata().cppparticledata,
                                                         pyglasstools.get_potential().cpppairpotential)
        solvers_list.append(self)

    def add_observables(self, observables):
        for name in observables:
            self.thermocalculator.addObservable(observables[name])

    def run(self):
        self.thermocalculator.compute()

    def update(self, frame_num):
        pyglasstools.update_sysdata(frame_num)
        self.thermocalculator.setSystemData(pyglasstools.get_sysdata().cppparticledata)
