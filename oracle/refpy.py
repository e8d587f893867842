"""TEST INFRASTRUCTURE — ctypes face of oracle/_ref/libadkref*.so (the unmodified reference solver).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


class _RefResult(ctypes.Structure):
    _fields_ = [
        ("found", ctypes.c_int32),
        ("count", ctypes.c_uint32),
        ("area", ctypes.c_double),
        ("hull_len", ctypes.c_uint32),
        ("diam_valid", ctypes.c_uint32),
        ("diam", ctypes.c_uint32 * 2),
        ("entries_alloc", ctypes.c_int64),
        ("entries_required", ctypes.c_int64),
        ("solve_ms", ctypes.c_double),
    ]


def ref_lib_path(variant="x87"):
    name = "libadkref.so" if variant == "x87" else "libadkref_f64.so"
    return os.path.join(_HERE, "_ref", name)


def ref_available(variant="x87"):
    return os.path.exists(ref_lib_path(variant))


class RefLib:
    """variant 'x87' = native long double build (the oracle); 'f64' = -mlong-double-64 build."""

    ALGO_EPPSTEIN, ALGO_ANTIPODAL, ALGO_ANTIPODAL_OPTIMIZED = 0, 1, 2

    def __init__(self, variant="x87"):
        self.lib = ctypes.CDLL(ref_lib_path(variant))
        self.lib.adkref_mant_dig.restype = ctypes.c_int
        self.mant_dig = self.lib.adkref_mant_dig()
        assert self.mant_dig == (64 if variant == "x87" else 53), self.mant_dig
        dp = ctypes.POINTER(ctypes.c_double)
        self.lib.adkref_solve.argtypes = [ctypes.c_int, dp, dp, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_double,
                                          ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_RefResult),
                                          ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32]
        self.lib.adkref_solve.restype = ctypes.c_int

    @staticmethod
    def _arr(v):
        return (ctypes.c_double * len(v))(*v)

    def solve(self, algo, x, y, kmax, amax, dmax, reconstruct=True, optimise=False):
        n = len(x)
        res = _RefResult()
        hull = (ctypes.c_uint32 * max(n, 1))()
        k = (2**64 - 1) if kmax is None or kmax >= 2**64 - 1 else int(kmax)
        rc = self.lib.adkref_solve(algo, self._arr(x), self._arr(y), n, k, float(amax), float(dmax),
                                   int(reconstruct), int(optimise), ctypes.byref(res), hull, n)
        assert rc == 0
        return {
            "found": int(res.found), "count": int(res.count), "area": float(res.area),
            "hull": [int(hull[i]) for i in range(res.hull_len)],
            "diam": [int(res.diam[0]), int(res.diam[1])] if res.diam_valid else None,
            "entries_alloc": int(res.entries_alloc), "entries_required": int(res.entries_required),
            "solve_ms": float(res.solve_ms),
        }

    def cw_order(self, x, y):
        n = len(x)
        out = (ctypes.c_int32 * (n * (n - 1)))()
        self.lib.adkref_cw_order(self._arr(x), self._arr(y), ctypes.c_uint32(n), out)
        return [[out[i * (n - 1) + p] for p in range(n - 1)] for i in range(n)]

    def count_tables(self, x, y):
        n = len(x)
        below = (ctypes.c_int32 * (n * n))()
        coll = (ctypes.c_int32 * (n * n))()
        rb = (ctypes.c_int32 * n)()
        self.lib.adkref_count_tables(self._arr(x), self._arr(y), ctypes.c_uint32(n), below, coll, rb)
        return list(below), list(coll), list(rb)

    def points_in_triangle(self, x, y, triples):
        n = len(x)
        m = len(triples)
        flat = (ctypes.c_uint32 * (3 * m))(*[v for t in triples for v in t])
        out = (ctypes.c_int32 * m)()
        self.lib.adkref_points_in_triangle(self._arr(x), self._arr(y), ctypes.c_uint32(n), flat, ctypes.c_uint32(m), out)
        return list(out)

    def antipodal_pairs(self, x, y, hull):
        """ForAllAntipodalPairs (Utils.h:57-109) over the polygon: list of (first, second) positions in call order."""
        n, m = len(x), len(hull)
        cap = 8 * m + 16
        out = (ctypes.c_uint32 * (2 * cap))()
        h = (ctypes.c_uint32 * m)(*hull)
        self.lib.adkref_antipodal_pairs.restype = ctypes.c_int
        k = self.lib.adkref_antipodal_pairs(self._arr(x), self._arr(y), ctypes.c_uint32(n), h, ctypes.c_uint32(m), out, ctypes.c_uint32(cap))
        assert k <= cap
        return [(int(out[2 * t]), int(out[2 * t + 1])) for t in range(k)]

    def compute_diameter(self, x, y, hull):
        n, m = len(x), len(hull)
        h = (ctypes.c_uint32 * m)(*hull)
        out = (ctypes.c_uint32 * 2)()
        self.lib.adkref_compute_diameter.restype = ctypes.c_int
        ok = self.lib.adkref_compute_diameter(self._arr(x), self._arr(y), ctypes.c_uint32(n), h, ctypes.c_uint32(m), out)
        return [int(out[0]), int(out[1])] if ok else None
