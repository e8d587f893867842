"""ctypes binding of the C ABI in include/adk.h (libadk.so).

This is plumbing for tests/ and bench.py; the product boundary is the C ABI itself and the C++ mirror of the
reference's namespace MT in host/.  There is no CPU fallback: without the built library or without a CUDA device
every entry point raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ADK_LIB_PATH") or os.path.join(_HERE, "libadk.so")     # ADK_LIB_PATH: development builds of the same library

ADK_OK, ADK_ERR_CUDA, ADK_ERR_ARG, ADK_ERR_NOMEM = 0, 1, 2, 3
FLAG_RECONSTRUCT, FLAG_COUNTERS = 1, 2
KMAX_INF = 2**64 - 1


class AdkResult(ctypes.Structure):
    _fields_ = [
        ("found", ctypes.c_int32), ("count", ctypes.c_uint32), ("area", ctypes.c_double),
        ("hull_len", ctypes.c_uint32), ("diam_valid", ctypes.c_uint32), ("diam", ctypes.c_uint32 * 2),
        ("entries_alloc", ctypes.c_int64), ("entries_required", ctypes.c_int64),
        ("cells_nominal", ctypes.c_uint64), ("cells_executed", ctypes.c_uint64),
        ("states", ctypes.c_uint64), ("pairs", ctypes.c_uint64),
        ("kernel_ms", ctypes.c_double), ("dp_ms", ctypes.c_double),
    ]


class AdkPartial(ctypes.Structure):
    _fields_ = [
        ("found", ctypes.c_int32), ("count", ctypes.c_uint32), ("area", ctypes.c_double),
        ("tie_key", ctypes.c_uint64), ("w0", ctypes.c_uint32), ("w1", ctypes.c_uint32),
    ]


class AdkStats(ctypes.Structure):
    _fields_ = [
        ("cells_nominal", ctypes.c_uint64), ("cells_executed", ctypes.c_uint64),
        ("states", ctypes.c_uint64), ("pairs", ctypes.c_uint64),
        ("kernel_ms", ctypes.c_double), ("dp_ms", ctypes.c_double),
        ("launches", ctypes.c_uint64), ("arena_bytes", ctypes.c_uint64),
    ]


class AdkInstance(ctypes.Structure):
    _fields_ = [
        ("x", ctypes.c_void_p), ("y", ctypes.c_void_p), ("n", ctypes.c_uint32), ("algo", ctypes.c_uint32),
        ("kmax", ctypes.c_uint64), ("amax", ctypes.c_double), ("dmax", ctypes.c_double),
        ("flags", ctypes.c_uint32), ("hull_cap", ctypes.c_uint32), ("hull", ctypes.c_void_p),
    ]


ALGO_AREA, ALGO_AREA_DIAMETER = 0, 1

EXPORTS = [
    "adk_create", "adk_destroy", "adk_last_error", "adk_set_memory_budget", "adk_set_stream", "adk_launch_count", "adk_set_reuse",
    "adk_area_finish_dev", "adk_area_diameter_finish_dev",
    "adk_solve_area", "adk_solve_area_diameter",
    "adk_area_partial", "adk_area_finish", "adk_area_diameter_partial", "adk_area_diameter_finish",
    "adk_merge", "adk_area_partial_dev", "adk_area_diameter_partial_dev",
    "adk_batch_assign", "adk_solve_batch",
    "adk_count_tables", "adk_points_in_triangle", "adk_cw_order", "adk_cw_order_noalloc", "adk_slope_order", "adk_lune_lists",
    "adk_solve_area_tiled",
    "adk_multi_create", "adk_multi_destroy", "adk_multi_last_error", "adk_multi_device_count", "adk_multi_set_reuse",
    "adk_multi_solve_area", "adk_multi_solve_area_diameter", "adk_multi_last_stats",
]


class AdkError(RuntimeError):
    pass


def load_library():
    if not os.path.exists(LIB_PATH):
        raise AdkError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    return ctypes.CDLL(LIB_PATH)


def _darr(v):
    return (ctypes.c_double * len(v))(*[float(t) for t in v])


def _k(kmax):
    return KMAX_INF if kmax is None or kmax >= KMAX_INF else int(kmax)


def make_instances(point_sets, algo, kmax, amax, dmax, reconstruct, counters):
    """point_sets: sequence of (n, 2) float64 arrays (NumPy; anything np.asarray takes).  The coordinates are split into
    contiguous SoA columns once; the C side borrows those buffers (no per-point marshalling)."""
    import numpy as np
    m = len(point_sets)
    inst = (AdkInstance * max(m, 1))()
    keep = []
    flags = (FLAG_RECONSTRUCT if reconstruct else 0) | (FLAG_COUNTERS if counters else 0)
    for i, pts in enumerate(point_sets):
        a = np.asarray(pts, dtype=np.float64).reshape(-1, 2)
        x = np.ascontiguousarray(a[:, 0]); y = np.ascontiguousarray(a[:, 1])
        hull = np.zeros(max(len(x) + 4, 4), dtype=np.uint32)
        keep.append((x, y, hull))
        inst[i].x = x.ctypes.data; inst[i].y = y.ctypes.data; inst[i].n = len(x); inst[i].algo = algo
        inst[i].kmax = _k(kmax); inst[i].amax = float(amax); inst[i].dmax = float(dmax)
        inst[i].flags = flags; inst[i].hull_cap = len(hull); inst[i].hull = hull.ctypes.data
    return inst, keep


def batch_assign(point_sets, nshards, algo=ALGO_AREA_DIAMETER):
    """adk_batch_assign: which rank solves which instance (host side, needs no device)."""
    lib = load_library()
    inst, _keep = make_instances(point_sets, algo, None, 0.0, 0.0, False, False)
    owner = (ctypes.c_uint32 * max(len(point_sets), 1))()
    lib.adk_batch_assign.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
    rc = lib.adk_batch_assign(ctypes.cast(inst, ctypes.c_void_p), len(point_sets), nshards, ctypes.cast(owner, ctypes.c_void_p))
    if rc != ADK_OK:
        raise AdkError(f"adk_batch_assign failed (status {rc})")
    return [int(owner[i]) for i in range(len(point_sets))]


class AdkMultiStats(ctypes.Structure):
    _fields_ = [
        ("ndev", ctypes.c_uint32), ("pad", ctypes.c_uint32),
        ("cells_nominal", ctypes.c_uint64), ("cells_executed", ctypes.c_uint64), ("states", ctypes.c_uint64),
        ("pairs", ctypes.c_uint64), ("launches", ctypes.c_uint64),
        ("max_kernel_ms", ctypes.c_double), ("max_dp_ms", ctypes.c_double), ("sum_dp_ms", ctypes.c_double),
        ("max_exchange_ms", ctypes.c_double),
    ]


class MultiSolver:
    """Several GPUs of one box behind one handle (adk_multi_*): one blocking call per solve; the shards, the NCCL all-gather of
    the 32-byte partials and the witness broadcast all happen inside libadk.so."""

    def __init__(self, devices=None):
        self.lib = load_library()
        L = self.lib
        L.adk_multi_last_error.restype = ctypes.c_char_p
        L.adk_multi_last_error.argtypes = [ctypes.c_void_p]
        L.adk_multi_create.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p, ctypes.c_int]
        L.adk_multi_destroy.argtypes = [ctypes.c_void_p]
        self.h = ctypes.c_void_p()
        if devices is None:
            rc = L.adk_multi_create(ctypes.byref(self.h), None, 0)
        else:
            arr = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
            rc = L.adk_multi_create(ctypes.byref(self.h), ctypes.cast(arr, ctypes.c_void_p), len(devices))
        if rc != ADK_OK:
            msg = L.adk_multi_last_error(None)
            raise AdkError(f"adk_multi_create failed ({rc}): {msg.decode() if msg else ''}")
        L.adk_multi_device_count.argtypes = [ctypes.c_void_p]
        self.ndev = int(L.adk_multi_device_count(self.h))

    def close(self):
        if self.h:
            self.lib.adk_multi_destroy(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != ADK_OK:
            msg = self.lib.adk_multi_last_error(self.h)
            raise AdkError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def set_reuse(self, enabled):
        self.lib.adk_multi_set_reuse.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self._check(self.lib.adk_multi_set_reuse(self.h, ctypes.c_int(1 if enabled else 0)), "adk_multi_set_reuse")

    def last_stats(self):
        st = AdkMultiStats()
        self.lib.adk_multi_last_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(AdkMultiStats)]
        self._check(self.lib.adk_multi_last_stats(self.h, ctypes.byref(st)), "adk_multi_last_stats")
        return {k: (float(getattr(st, k)) if k.endswith("_ms") else int(getattr(st, k))) for k, _ in AdkMultiStats._fields_ if k != "pad"}

    def solve_area_host(self, px, py, n, kmax, amax, reconstruct=True, counters=False):
        """Host pointers (e.g. pinned torch tensors' data_ptr()) in, dict out."""
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n, 1))()
        flags = (FLAG_RECONSTRUCT if reconstruct else 0) | (FLAG_COUNTERS if counters else 0)
        self.lib.adk_multi_solve_area.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64,
                                                  ctypes.c_double, ctypes.c_uint32, ctypes.POINTER(AdkResult), ctypes.c_void_p]
        self._check(self.lib.adk_multi_solve_area(self.h, ctypes.c_void_p(px), ctypes.c_void_p(py), n, _k(kmax), float(amax), flags,
                                                  ctypes.byref(res), ctypes.cast(hull, ctypes.c_void_p)), "adk_multi_solve_area")
        return Solver._pack(res, hull)

    def solve_area_diameter_host(self, px, py, n, kmax, amax, dmax, reconstruct=True, counters=False):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n, 1))()
        flags = (FLAG_RECONSTRUCT if reconstruct else 0) | (FLAG_COUNTERS if counters else 0)
        self.lib.adk_multi_solve_area_diameter.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64,
                                                           ctypes.c_double, ctypes.c_double, ctypes.c_uint32, ctypes.POINTER(AdkResult),
                                                           ctypes.c_void_p]
        self._check(self.lib.adk_multi_solve_area_diameter(self.h, ctypes.c_void_p(px), ctypes.c_void_p(py), n, _k(kmax), float(amax),
                                                           float(dmax), flags, ctypes.byref(res), ctypes.cast(hull, ctypes.c_void_p)),
                    "adk_multi_solve_area_diameter")
        return Solver._pack(res, hull)

    def solve_area(self, x, y, kmax=None, amax=float("inf"), reconstruct=True, counters=False):
        ax, ay = _darr(x), _darr(y)
        return self.solve_area_host(ctypes.addressof(ax), ctypes.addressof(ay), len(x), kmax, amax, reconstruct, counters)

    def solve_area_diameter(self, x, y, kmax=None, amax=float("inf"), dmax=float("inf"), reconstruct=True, counters=False):
        ax, ay = _darr(x), _darr(y)
        return self.solve_area_diameter_host(ctypes.addressof(ax), ctypes.addressof(ay), len(x), kmax, amax, dmax, reconstruct, counters)


class Solver:
    """One context = one GPU.  Mirrors the reference's call shapes; returns plain dicts."""

    def __init__(self, device=-1):
        self.lib = load_library()
        L = self.lib
        L.adk_last_error.restype = ctypes.c_char_p
        L.adk_last_error.argtypes = [ctypes.c_void_p]
        L.adk_create.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int]
        L.adk_destroy.argtypes = [ctypes.c_void_p]
        L.adk_set_memory_budget.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        self.ctx = ctypes.c_void_p()
        rc = L.adk_create(ctypes.byref(self.ctx), int(device))
        if rc != ADK_OK:
            msg = L.adk_last_error(None)
            raise AdkError(f"adk_create failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if self.ctx:
            self.lib.adk_destroy(self.ctx)
            self.ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != ADK_OK:
            msg = self.lib.adk_last_error(self.ctx)
            raise AdkError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def set_stream(self, cuda_stream):
        """Run on the caller's stream (an integer cudaStream_t handle, 0 = legacy default stream)."""
        self.lib.adk_set_stream.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        self._check(self.lib.adk_set_stream(self.ctx, ctypes.c_void_p(int(cuda_stream))), "adk_set_stream")

    def launch_count(self):
        self.lib.adk_launch_count.restype = ctypes.c_uint64
        self.lib.adk_launch_count.argtypes = [ctypes.c_void_p]
        return int(self.lib.adk_launch_count(self.ctx))

    def set_reuse(self, enabled):
        """False: every solve recomputes the count tables and angular orders (what the reference does on every call)."""
        self.lib.adk_set_reuse.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self._check(self.lib.adk_set_reuse(self.ctx, ctypes.c_int(1 if enabled else 0)), "adk_set_reuse")

    def set_memory_budget(self, nbytes):
        self._check(self.lib.adk_set_memory_budget(self.ctx, ctypes.c_uint64(int(nbytes))), "adk_set_memory_budget")

    @staticmethod
    def _pack(res, hull):
        return {
            "found": int(res.found), "count": int(res.count), "area": float(res.area),
            "hull": [int(hull[i]) for i in range(res.hull_len)],
            "diam": [int(res.diam[0]), int(res.diam[1])] if res.diam_valid else None,
            "entries_alloc": int(res.entries_alloc), "entries_required": int(res.entries_required),
            "cells_nominal": int(res.cells_nominal), "cells_executed": int(res.cells_executed),
            "states": int(res.states), "pairs": int(res.pairs),
            "kernel_ms": float(res.kernel_ms), "dp_ms": float(res.dp_ms),
        }

    @staticmethod
    def _part(p):
        return {"found": int(p.found), "count": int(p.count), "area": float(p.area), "tie_key": int(p.tie_key),
                "w0": int(p.w0), "w1": int(p.w1)}

    @staticmethod
    def _unpart(d):
        return AdkPartial(d["found"], d["count"], d["area"], d["tie_key"], d["w0"], d["w1"])

    @staticmethod
    def _stats(s):
        return {k: (float(getattr(s, k)) if k.endswith("_ms") else int(getattr(s, k))) for k, _ in AdkStats._fields_}

    def solve_area_tiled(self, points, patch_size=3.0, step_size=3 * 0.66, kmax=None, amax=4.0):
        """run_area_only_tiled (Area-Selector) in one C call, windows generated on the device (adk_solve_area_tiled).
        Returns {found, area, count, hull (indices into `points`), diam, patches, solved}."""
        n = len(points)
        ax, ay = _darr([p[0] for p in points]), _darr([p[1] for p in points])
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n, 1))()
        st = (ctypes.c_uint64 * 2)()
        self.lib.adk_solve_area_tiled.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_double,
                                                  ctypes.c_double, ctypes.c_uint64, ctypes.c_double, ctypes.POINTER(AdkResult),
                                                  ctypes.c_void_p, ctypes.c_void_p]
        self._check(self.lib.adk_solve_area_tiled(self.ctx, ctypes.cast(ax, ctypes.c_void_p), ctypes.cast(ay, ctypes.c_void_p), n,
                                                  float(patch_size), float(step_size), _k(kmax), float(amax), ctypes.byref(res),
                                                  ctypes.cast(hull, ctypes.c_void_p), ctypes.cast(st, ctypes.c_void_p)), "adk_solve_area_tiled")
        out = self._pack(res, hull)
        out["patches"], out["solved"] = int(st[0]), int(st[1])
        return out

    # ---- whole solves -------------------------------------------------------------------------------------------
    def solve_area(self, x, y, kmax=None, amax=float("inf"), reconstruct=True, counters=False):
        n = len(x)
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        flags = (FLAG_RECONSTRUCT if reconstruct else 0) | (FLAG_COUNTERS if counters else 0)
        rc = self.lib.adk_solve_area(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint64(_k(kmax)),
                                     ctypes.c_double(amax), ctypes.c_uint32(flags), ctypes.byref(res), hull)
        self._check(rc, "adk_solve_area")
        return self._pack(res, hull)

    def solve_area_diameter(self, x, y, kmax=None, amax=float("inf"), dmax=float("inf"), reconstruct=True, counters=False):
        n = len(x)
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        flags = (FLAG_RECONSTRUCT if reconstruct else 0) | (FLAG_COUNTERS if counters else 0)
        rc = self.lib.adk_solve_area_diameter(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint64(_k(kmax)),
                                              ctypes.c_double(amax), ctypes.c_double(dmax), ctypes.c_uint32(flags),
                                              ctypes.byref(res), hull)
        self._check(rc, "adk_solve_area_diameter")
        return self._pack(res, hull)

    # ---- batches of independent instances (BASELINE config 5) -----------------------------------------
    _instances = staticmethod(lambda *a: make_instances(*a))

    def solve_batch(self, point_sets, algo=ALGO_AREA_DIAMETER, kmax=None, amax=float("inf"), dmax=float("inf"),
                    reconstruct=True, counters=False, shard=0, nshards=1):
        """One call for a list of independent point sets.  Returns one dict per instance (None for instances owned by
        another shard)."""
        m = len(point_sets)
        inst, keep = self._instances(point_sets, algo, kmax, amax, dmax, reconstruct, counters)
        res = (AdkResult * max(m, 1))()
        self.lib.adk_solve_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]
        rc = self.lib.adk_solve_batch(self.ctx, ctypes.cast(inst, ctypes.c_void_p), m, shard, nshards, ctypes.cast(res, ctypes.c_void_p))
        self._check(rc, "adk_solve_batch")
        return [None if res[i].found < 0 else self._pack(res[i], keep[i][2]) for i in range(m)]

    # ---- sharded solves ------------------------------------------------------------------------------------------
    def area_partial(self, x, y, kmax, amax, shard, nshards):
        p, s = AdkPartial(), AdkStats()
        rc = self.lib.adk_area_partial(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(len(x)), ctypes.c_uint64(_k(kmax)),
                                       ctypes.c_double(amax), ctypes.c_uint32(shard), ctypes.c_uint32(nshards),
                                       ctypes.byref(p), ctypes.byref(s))
        self._check(rc, "adk_area_partial")
        return self._part(p), self._stats(s)

    def area_diameter_partial(self, x, y, kmax, amax, dmax, shard, nshards):
        p, s = AdkPartial(), AdkStats()
        rc = self.lib.adk_area_diameter_partial(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(len(x)),
                                                ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                                ctypes.c_uint32(shard), ctypes.c_uint32(nshards), ctypes.byref(p), ctypes.byref(s))
        self._check(rc, "adk_area_diameter_partial")
        return self._part(p), self._stats(s)

    def area_partial_dev(self, d_x, d_y, n, kmax, amax, shard=0, nshards=1):
        p, s = AdkPartial(), AdkStats()
        rc = self.lib.adk_area_partial_dev(self.ctx, ctypes.c_void_p(d_x), ctypes.c_void_p(d_y), ctypes.c_uint32(n),
                                           ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_uint32(shard),
                                           ctypes.c_uint32(nshards), ctypes.byref(p), ctypes.byref(s))
        self._check(rc, "adk_area_partial_dev")
        return self._part(p), self._stats(s)

    def area_diameter_partial_dev(self, d_x, d_y, n, kmax, amax, dmax, shard=0, nshards=1):
        p, s = AdkPartial(), AdkStats()
        rc = self.lib.adk_area_diameter_partial_dev(self.ctx, ctypes.c_void_p(d_x), ctypes.c_void_p(d_y), ctypes.c_uint32(n),
                                                    ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                                    ctypes.c_uint32(shard), ctypes.c_uint32(nshards), ctypes.byref(p),
                                                    ctypes.byref(s))
        self._check(rc, "adk_area_diameter_partial_dev")
        return self._part(p), self._stats(s)

    def merge(self, parts):
        arr = (AdkPartial * len(parts))(*[self._unpart(p) for p in parts])
        out = AdkPartial()
        self._check(self.lib.adk_merge(arr, ctypes.c_uint32(len(parts)), ctypes.byref(out)), "adk_merge")
        return self._part(out)

    def area_finish(self, x, y, kmax, amax, best, reconstruct=True):
        n = len(x)
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        b = self._unpart(best)
        rc = self.lib.adk_area_finish(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint64(_k(kmax)),
                                      ctypes.c_double(amax), ctypes.byref(b), ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0),
                                      ctypes.byref(res), hull)
        self._check(rc, "adk_area_finish")
        return self._pack(res, hull)

    def area_diameter_finish(self, x, y, kmax, amax, dmax, best, reconstruct=True):
        n = len(x)
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        b = self._unpart(best)
        rc = self.lib.adk_area_diameter_finish(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint64(_k(kmax)),
                                               ctypes.c_double(amax), ctypes.c_double(dmax), ctypes.byref(b),
                                               ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0), ctypes.byref(res), hull)
        self._check(rc, "adk_area_diameter_finish")
        return self._pack(res, hull)

    def area_diameter_finish_dev(self, d_x, d_y, n, kmax, amax, dmax, best, reconstruct=True):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        b = self._unpart(best)
        rc = self.lib.adk_area_diameter_finish_dev(self.ctx, ctypes.c_void_p(d_x), ctypes.c_void_p(d_y), ctypes.c_uint32(n),
                                                   ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                                   ctypes.byref(b), ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0),
                                                   ctypes.byref(res), hull)
        self._check(rc, "adk_area_diameter_finish_dev")
        return self._pack(res, hull)

    def area_finish_dev(self, d_x, d_y, n, kmax, amax, best, reconstruct=True):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        b = self._unpart(best)
        rc = self.lib.adk_area_finish_dev(self.ctx, ctypes.c_void_p(d_x), ctypes.c_void_p(d_y), ctypes.c_uint32(n),
                                          ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.byref(b),
                                          ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0), ctypes.byref(res), hull)
        self._check(rc, "adk_area_finish_dev")
        return self._pack(res, hull)

    # ---- raw host-pointer variants (bench.py: pinned host buffers go straight through the C ABI) ------------------
    def solve_area_diameter_host(self, px, py, n, kmax, amax, dmax, reconstruct=True):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        rc = self.lib.adk_solve_area_diameter(self.ctx, ctypes.c_void_p(px), ctypes.c_void_p(py), ctypes.c_uint32(n),
                                              ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                              ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0), ctypes.byref(res), hull)
        self._check(rc, "adk_solve_area_diameter")
        return self._pack(res, hull)

    def solve_area_host(self, px, py, n, kmax, amax, reconstruct=True):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        rc = self.lib.adk_solve_area(self.ctx, ctypes.c_void_p(px), ctypes.c_void_p(py), ctypes.c_uint32(n),
                                     ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax),
                                     ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0), ctypes.byref(res), hull)
        self._check(rc, "adk_solve_area")
        return self._pack(res, hull)

    def area_diameter_partial_host(self, px, py, n, kmax, amax, dmax, shard, nshards):
        p, s = AdkPartial(), AdkStats()
        rc = self.lib.adk_area_diameter_partial(self.ctx, ctypes.c_void_p(px), ctypes.c_void_p(py), ctypes.c_uint32(n),
                                                ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                                ctypes.c_uint32(shard), ctypes.c_uint32(nshards), ctypes.byref(p), ctypes.byref(s))
        self._check(rc, "adk_area_diameter_partial")
        return self._part(p), self._stats(s)

    def area_diameter_finish_host(self, px, py, n, kmax, amax, dmax, best, reconstruct=True):
        res = AdkResult()
        hull = (ctypes.c_uint32 * max(n + 4, 4))()
        b = self._unpart(best)
        rc = self.lib.adk_area_diameter_finish(self.ctx, ctypes.c_void_p(px), ctypes.c_void_p(py), ctypes.c_uint32(n),
                                               ctypes.c_uint64(_k(kmax)), ctypes.c_double(amax), ctypes.c_double(dmax),
                                               ctypes.byref(b), ctypes.c_uint32(FLAG_RECONSTRUCT if reconstruct else 0),
                                               ctypes.byref(res), hull)
        self._check(rc, "adk_area_diameter_finish")
        return self._pack(res, hull)

    # ---- precompute stages ---------------------------------------------------------------------------------------
    def count_tables(self, x, y):
        n = len(x)
        below = (ctypes.c_int32 * (n * n))()
        coll = (ctypes.c_int32 * (n * n))()
        rb = (ctypes.c_int32 * n)()
        self._check(self.lib.adk_count_tables(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), below, coll, rb), "adk_count_tables")
        return list(below), list(coll), list(rb)

    def points_in_triangle(self, x, y, triples):
        m = len(triples)
        flat = (ctypes.c_uint32 * (3 * m))(*[v for t in triples for v in t])
        out = (ctypes.c_int32 * m)()
        self._check(self.lib.adk_points_in_triangle(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(len(x)), flat,
                                                    ctypes.c_uint32(m), out), "adk_points_in_triangle")
        return list(out)

    def cw_order(self, x, y):
        n = len(x)
        out = (ctypes.c_int32 * (n * (n - 1)))()
        self._check(self.lib.adk_cw_order(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), out), "adk_cw_order")
        return [[out[i * (n - 1) + p] for p in range(n - 1)] for i in range(n)]

    def slope_order(self, x, y, i, j):
        n = len(x)
        out = (ctypes.c_int32 * (n + 4))()
        ln = ctypes.c_uint32()
        self._check(self.lib.adk_slope_order(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint32(i),
                                             ctypes.c_uint32(j), out, ctypes.byref(ln)), "adk_slope_order")
        return [out[t] for t in range(ln.value)]

    def lune_lists(self, x, y, s, t):
        n = len(x)
        left = (ctypes.c_int32 * (n + 4))()
        right = (ctypes.c_int32 * (n + 4))()
        nl, nr = ctypes.c_uint32(), ctypes.c_uint32()
        self._check(self.lib.adk_lune_lists(self.ctx, _darr(x), _darr(y), ctypes.c_uint32(n), ctypes.c_uint32(s),
                                            ctypes.c_uint32(t), left, ctypes.byref(nl), right, ctypes.byref(nr)), "adk_lune_lists")
        return [left[i] for i in range(nl.value)], [right[i] for i in range(nr.value)]
