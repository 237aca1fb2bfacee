"""vpic_b200 -- B200-native (sm_100a) implementation of VPIC's per-step PIC hot
path behind the reference's own species / sf_interface / field_advance entry
points.  The product is vpic_b200/libvpic_b200.so (C ABI, include/vpic_b200/);
this package is the thin Python harness used by the tests and bench.py."""
from .lib import load_b200, b200_backend, Backend, B200_LIB  # noqa: F401
from .driver import Simulation, B200Host                      # noqa: F401
