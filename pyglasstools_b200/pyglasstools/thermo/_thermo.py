"""thermo/module-thermo.cpp:7-45 names, re-exported from the host extension."""
from pyglasstools._pgt_host import (GlobalBornTensor2D, GlobalBornTensor3D, GlobalElasticVirialStress2D,  # noqa: F401
                                    GlobalElasticVirialStress3D, GlobalKineticStress2D, GlobalKineticStress3D,
                                    GlobalPotentialEnergy2D, GlobalPotentialEnergy3D, GlobalVirialStress2D,
                                    GlobalVirialStress3D, ThermoCalculator)
