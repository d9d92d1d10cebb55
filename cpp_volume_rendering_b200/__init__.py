"""cpp_volume_rendering_b200 — B200-native structured single-pass volume ray casting.

The product is libvrb200.so (hand-written sm_100a CUDA behind the C ABI in include/vrb200.h) and the
C++ plugins in host/ that mirror the reference's BaseVolumeRenderer interface.  This Python package
is only the ctypes plumbing tests and bench.py use to reach the C ABI.
"""
from . import _capi, inputs, metrics  # noqa: F401
