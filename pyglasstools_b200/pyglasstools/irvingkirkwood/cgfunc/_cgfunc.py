"""irvingkirkwood/cgfunc/module-cgfunc.cpp:12-30 names, re-exported from the host extension."""
from pyglasstools._pgt_host import (CGFuncMollifierAdaptive, CGFuncMollifierFixedPoint, CGFuncOcticAdaptive,  # noqa: F401
                                    CGFuncOcticFixedPoint, CGFuncRectAdaptive, CGFuncRectFixedPoint,
                                    CoarseGrainFunction)
