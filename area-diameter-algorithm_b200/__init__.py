"""B200-native (sm_100a) solver for the two hot paths of gianmarcopicarella/area-diameter-algorithm.

Layout: csrc/ = CUDA kernels + the C ABI (include/adk.h), host/ = C++ mirror of the reference's namespace MT,
capi.py = ctypes plumbing used by tests/ and bench.py.
"""
from .capi import (Solver, MultiSolver, AdkError, load_library, LIB_PATH, EXPORTS, KMAX_INF, batch_assign,  # noqa: F401
                   ALGO_AREA, ALGO_AREA_DIAMETER)
