"""Private module name of the reference (module-pyglasstools.cpp:59-71), re-exported from the single host extension."""
from pyglasstools._pgt_host import (Calculator, Communicator, Manager, MPILogFile, Observable, ParticleSystem,  # noqa: F401
                                    SimBox, array_equal_fast, device_close, device_launches, format_field_frame, format_field_frame_bytes, write_field_frame,
                                    ostream_double,
                                    repr_double)
