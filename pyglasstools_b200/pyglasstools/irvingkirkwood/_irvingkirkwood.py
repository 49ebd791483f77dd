"""irvingkirkwood/module-irvingkirkwood.cpp:8-32 names, re-exported from the host extension."""
from pyglasstools._pgt_host import (ElasticVirialStressField2D, ElasticVirialStressField3D, IrvingKirkwood,  # noqa: F401
                                    KineticStressField2D, KineticStressField3D, VirialStressField2D,
                                    VirialStressField3D)
