"""adk_cw_order_noalloc: the clockwise-order stage with its scratch in a static device array (no cudaMalloc inside the call).

No counterpart in the reference (include/adk.h).  The file is
named so that pytest collects it last: a failure here must not hide the parity tests.
It passes alone in a fresh process and after the other tests (it creates its own context and its own device buffers).
"""
import ctypes
import os
import random
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

MAX_N = 512          # ADK_NOALLOC_MAX_POINTS (include/adk.h)


def _noalloc_order(solver, torch, x, y):
    """Calls the C-ABI entry point on caller-owned device buffers; returns the n lists of n-1 indices."""
    n = len(x)
    dev = torch.device("cuda", 0)
    d_x = torch.tensor(x, dtype=torch.float64, device=dev)
    d_y = torch.tensor(y, dtype=torch.float64, device=dev)
    d_order = torch.full((n * (n - 1),), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    f = solver.lib.adk_cw_order_noalloc
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
    f.restype = ctypes.c_int
    rc = f(solver.ctx, ctypes.c_void_p(d_x.data_ptr()), ctypes.c_void_p(d_y.data_ptr()), n, ctypes.c_void_p(d_order.data_ptr()))
    torch.cuda.synchronize()            # the call queues its kernels on the context's stream and does not synchronise
    return rc, d_order.cpu().tolist()


def test_no_alloc_clockwise_order_equals_the_allocating_entry_point_twice_in_a_row():
    import torch
    import __graft_entry__ as ge
    ge.build()
    pkg = ge.load_package()
    solver = pkg.Solver(0)
    try:
        rng = random.Random(31)
        # two calls in a row on the shared static scratch: a large set first, then a smaller one whose tables overlap the
        # stale contents of the first call (index work: the lists must be identical, not merely close)
        for n in (MAX_N, 137):
            x = [rng.uniform(0, 20) for _ in range(n)]
            y = [rng.uniform(0, 20) for _ in range(n)]
            want = solver.cw_order(x, y)                                   # the existing entry point (context buffers)
            rc, flat = _noalloc_order(solver, torch, x, y)
            assert rc == 0, solver.lib.adk_last_error(solver.ctx)
            got = [flat[i * (n - 1):(i + 1) * (n - 1)] for i in range(n)]
            assert got == want, n
        # near ties exercise the exact-comparator fix-up through the same scratch
        x = [10.0] + [10.0 + r * c for c in (1.0, 0.6, -0.8) for r in (1.0, 2.0, 3.5)]
        y = [10.0] + [10.0 + r * s for s in (0.0, 0.8, 0.6) for r in (1.0, 2.0, 3.5)]
        rc, flat = _noalloc_order(solver, torch, x, y)
        n = len(x)
        assert rc == 0 and [flat[i * (n - 1):(i + 1) * (n - 1)] for i in range(n)] == solver.cw_order(x, y)
        # beyond the static scratch: a clear error, nothing launched
        n = MAX_N + 1
        rc, flat = _noalloc_order(solver, torch, [float(i) for i in range(n)], [float(i * i % 97) for i in range(n)])
        assert rc == 2                                                      # ADK_ERR_ARG
        msg = solver.lib.adk_last_error(solver.ctx).decode()
        assert "ADK_NOALLOC_MAX_POINTS" in msg and str(n) in msg, msg
        assert all(v == -1 for v in flat[:1000])                            # the output was not touched
    finally:
        solver.close()
