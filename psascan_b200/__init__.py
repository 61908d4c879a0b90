"""psascan_b200 -- B200-native implementation of pSAscan's data-parallel hot path.

The product is libpsascan_cuda.so (hand-written sm_100a CUDA behind the C ABI declared in
include/psascan_cuda.h); this package is the thin ctypes binding used by tests and bench.py.
There is no CPU fallback: every compute call needs a CUDA device.
"""
from .lib import (Context, PscError, load_library, library_path, EXPORTS)  # noqa: F401
from . import textgen  # noqa: F401
