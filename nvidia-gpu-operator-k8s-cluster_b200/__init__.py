"""b200-node-validator: the B200-native (sm_100a) cuda-vectoradd validation workload.

The product is native: csrc/ (CUDA kernels + the C ABI of include/b200validate.h), host/ (the
C++ `vectorAdd` and `nvidia-validator` executables) and go/ (the cgo driver source).  This Python
package is only the ctypes binding the tests and bench.py call the C ABI through, plus the
rank-sharding helper for the per-GPU fan-out.  It has no CPU fallback.
"""
from . import capi  # noqa: F401
from .capi import B200VError  # noqa: F401
from . import fanout  # noqa: F401
