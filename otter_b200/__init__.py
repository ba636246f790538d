"""otter_b200 — B200-native (sm_100a) hot path of holstegelab/otter behind a C ABI.

The Python layer here is only a ctypes binding of include/otter_b200.h for tests and bench.py;
the product is otter_b200/lib/libotter_b200.so (CUDA) plus the C++17 host mirror under
otter_b200/host/.  There is no CPU fallback: without the built library or without a GPU the
compute calls raise.
"""
from .capi import Context, OtgpuError, lib, LIB_PATH  # noqa: F401
