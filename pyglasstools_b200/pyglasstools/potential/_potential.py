"""potential/module-potential.cpp:13-34 names, re-exported from the host extension."""
from pyglasstools._pgt_host import (PairPotential, PairPotentialForceShiftedLJ, PairPotentialLJ,  # noqa: F401
                                    PairPotentialPoly10, PairPotentialPoly12, PairPotentialPoly18,
                                    PairPotentialPoly106, PairPotentialPolyLJ, PairPotentialPolyYukawa)
