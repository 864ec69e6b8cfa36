"""cfd-nn_b200 — B200-native backend for the fractional-step hot path of comp-physics/cfd-nn.

The product is cfd-nn_b200/libcfdnn_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/cfdnn_b200.h).  The C++ host-side mirror of the reference interface (RANSSolver /
MultigridPoissonSolver / TurbulenceModel shims) is include/cfdnn_b200_shim.hpp; this package holds
the ctypes binding (`_capi`) and the Python mirror of the same call sequence (`backend`) that
bench.py, smoke() and the tests drive.  The directory name carries a hyphen, so it is imported from
its path (see `load_package` in bench.py / __graft_entry__.py), under the module name `cfdnn_b200`.
"""
from . import _capi  # noqa: F401
from . import backend  # noqa: F401
from ._capi import load, DEFAULT_LIB, EXPORTS  # noqa: F401
from .backend import Backend  # noqa: F401
