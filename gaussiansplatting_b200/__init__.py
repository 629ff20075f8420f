"""B200-native tiled Gaussian-splatting hot path behind the reference's operator interface.

The product is libgs_b200.so (hand-written sm_100a CUDA kernels + the C ABI of include/gs_abi.h); `capi` is the
ctypes binding used by tests and bench, `host/` the C++ mirror of the reference's classes.  There is no CPU
fallback: anything that computes raises GsError when the CUDA library is missing.
"""
from . import layouts, synth  # noqa: F401
from .capi import GsError, LIB_PATH  # noqa: F401
