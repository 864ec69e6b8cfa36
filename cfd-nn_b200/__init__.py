"""cfd-nn_b200 — B200-native backend for the fractional-step hot path of comp-physics/cfd-nn.

The product is cfd-nn_b200/libcfdnn_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/cfdnn_b200.h).  This package only holds the ctypes binding used by tests/ and bench.py;
the C++ host-side mirror of the reference interface is cfd-nn_b200/host/cfdnn_b200.hpp.
"""
from . import _capi  # noqa: F401
from ._capi import load, DEFAULT_LIB, EXPORTS  # noqa: F401
