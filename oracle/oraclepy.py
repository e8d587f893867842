"""TEST INFRASTRUCTURE — ctypes face of oracle/libadkoracle.so (the C restatement in adk_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libadkoracle.so")


class _Res(ctypes.Structure):
    _fields_ = [
        ("found", ctypes.c_int32), ("count", ctypes.c_uint32), ("area", ctypes.c_double),
        ("hull_len", ctypes.c_uint32), ("diam_valid", ctypes.c_uint32), ("diam", ctypes.c_uint32 * 2),
        ("entries_alloc", ctypes.c_int64), ("entries_required", ctypes.c_int64),
        ("cells", ctypes.c_uint64), ("states", ctypes.c_uint64), ("pairs", ctypes.c_uint64),
        ("solve_ms", ctypes.c_double),
    ]


def build_oracle():
    src = os.path.join(_HERE, "adk_oracle.c")
    if (not os.path.exists(_LIB)) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libadkoracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


def _arr(v):
    return (ctypes.c_double * len(v))(*[float(t) for t in v])


class Oracle:
    KINF = 2**64 - 1

    def __init__(self):
        self.lib = ctypes.CDLL(build_oracle())

    @staticmethod
    def _pack(res, hull):
        return {
            "found": int(res.found), "count": int(res.count), "area": float(res.area),
            "hull": [int(hull[i]) for i in range(res.hull_len)],
            "diam": [int(res.diam[0]), int(res.diam[1])] if res.diam_valid else None,
            "entries_alloc": int(res.entries_alloc), "entries_required": int(res.entries_required),
            "cells": int(res.cells), "states": int(res.states), "pairs": int(res.pairs),
            "solve_ms": float(res.solve_ms),
        }

    @classmethod
    def _k(cls, kmax):
        return cls.KINF if kmax is None or kmax >= cls.KINF else int(kmax)

    def solve_area(self, x, y, kmax=None, amax=float("inf"), reconstruct=True):
        n = len(x)
        res = _Res()
        hull = (ctypes.c_uint32 * max(n + 2, 1))()
        rc = self.lib.adko_solve_area(_arr(x), _arr(y), ctypes.c_uint32(n), ctypes.c_uint64(self._k(kmax)),
                                      ctypes.c_double(amax), ctypes.c_int(int(reconstruct)), ctypes.byref(res), hull)
        assert rc == 0, rc
        return self._pack(res, hull)

    def solve_area_diameter(self, x, y, kmax=None, amax=float("inf"), dmax=float("inf"), reconstruct=True):
        n = len(x)
        res = _Res()
        hull = (ctypes.c_uint32 * max(n + 2, 1))()
        rc = self.lib.adko_solve_area_diameter(_arr(x), _arr(y), ctypes.c_uint32(n), ctypes.c_uint64(self._k(kmax)),
                                               ctypes.c_double(amax), ctypes.c_double(dmax),
                                               ctypes.c_int(int(reconstruct)), ctypes.byref(res), hull)
        assert rc == 0, rc
        return self._pack(res, hull)

    def solve_area_diameter_shard(self, x, y, kmax, amax, dmax, shard, nshards):
        """Every nshards-th candidate pair only (no reconstruction); merge with merge_shards."""
        res = _Res()
        hull = (ctypes.c_uint32 * 1)()
        rc = self.lib.adko_solve_area_diameter_shard(_arr(x), _arr(y), ctypes.c_uint32(len(x)), ctypes.c_uint64(self._k(kmax)),
                                                     ctypes.c_double(amax), ctypes.c_double(dmax), ctypes.c_uint32(shard),
                                                     ctypes.c_uint32(nshards), ctypes.byref(res))
        assert rc == 0, rc
        return self._pack(res, hull)

    def solve_area_shard(self, x, y, kmax, amax, reconstruct, shard, nshards):
        """Every nshards-th anchor of the bottom-up order only; merge with merge_area_shards."""
        n = len(x)
        res = _Res()
        hull = (ctypes.c_uint32 * max(n + 2, 1))()
        pos = ctypes.c_uint32(0)
        rc = self.lib.adko_solve_area_shard(_arr(x), _arr(y), ctypes.c_uint32(n), ctypes.c_uint64(self._k(kmax)), ctypes.c_double(amax),
                                            ctypes.c_int(int(reconstruct)), ctypes.c_uint32(shard), ctypes.c_uint32(nshards),
                                            ctypes.byref(res), hull, ctypes.byref(pos))
        assert rc == 0, rc
        out = self._pack(res, hull)
        out["anchor_pos"] = int(pos.value)
        return out

    @staticmethod
    def merge_area_shards(parts):
        """Best of the shards: count desc, area asc, anchor position asc (Eppstein.cpp:219 replaces the running best only by a strictly
        better one, so the first anchor of equals stays); cells and entries_required add up."""
        best = None
        for p in parts:
            if p["found"] and (best is None or (-p["count"], p["area"], p["anchor_pos"]) < (-best["count"], best["area"], best["anchor_pos"])):
                best = p
        out = dict(best) if best else {"found": 0, "count": 0, "area": 0.0, "hull": [], "diam": None}
        out["cells"] = sum(p["cells"] for p in parts)
        out["entries_required"] = sum(p["entries_required"] for p in parts)
        return out

    @staticmethod
    def merge_shards(parts):
        """Best of the shards: count desc, area asc, pair in enumeration order asc (the strict comparisons of
        AntipodalOptimized.cpp:512-517 keep the first of equals); cells and states add up."""
        best = None
        for p in parts:
            if p["found"] and (best is None or (-p["count"], p["area"], p["diam"]) < (-best["count"], best["area"], best["diam"])):
                best = p
        out = dict(best) if best else {"found": 0, "count": 0, "area": 0.0, "diam": None}
        out["cells"] = sum(p["cells"] for p in parts); out["states"] = sum(p["states"] for p in parts)
        out["pairs"] = parts[0]["pairs"] if parts else 0
        return out

    def cw_order(self, x, y):
        n = len(x)
        out = (ctypes.c_int32 * (n * (n - 1)))()
        self.lib.adko_cw_order(_arr(x), _arr(y), ctypes.c_uint32(n), out)
        return [[out[i * (n - 1) + p] for p in range(n - 1)] for i in range(n)]

    def _tables(self, fn, x, y):
        n = len(x)
        below = (ctypes.c_int32 * (n * n))()
        coll = (ctypes.c_int32 * (n * n))()
        rb = (ctypes.c_int32 * n)()
        fn(_arr(x), _arr(y), ctypes.c_uint32(n), below, coll, rb)
        return list(below), list(coll), list(rb)

    def count_tables(self, x, y):
        return self._tables(self.lib.adko_count_tables, x, y)

    def count_tables_bruteforce(self, x, y):
        return self._tables(self.lib.adko_count_tables_bruteforce, x, y)

    def points_in_triangle(self, x, y, triples):
        m = len(triples)
        flat = (ctypes.c_uint32 * (3 * m))(*[v for t in triples for v in t])
        out = (ctypes.c_int32 * m)()
        self.lib.adko_points_in_triangle(_arr(x), _arr(y), ctypes.c_uint32(len(x)), flat, ctypes.c_uint32(m), out)
        return list(out)

    def slope_order(self, x, y, i, j):
        n = len(x)
        out = (ctypes.c_int32 * n)()
        c = self.lib.adko_slope_order(_arr(x), _arr(y), ctypes.c_uint32(n), ctypes.c_uint32(i), ctypes.c_uint32(j), out)
        assert c >= 0
        return [out[t] for t in range(c)]

    def lune_lists(self, x, y, s, t):
        n = len(x)
        left = (ctypes.c_int32 * (n + 2))()
        right = (ctypes.c_int32 * (n + 2))()
        nl = ctypes.c_uint32()
        nr = ctypes.c_uint32()
        self.lib.adko_lune_lists(_arr(x), _arr(y), ctypes.c_uint32(n), ctypes.c_uint32(s), ctypes.c_uint32(t),
                                 left, ctypes.byref(nl), right, ctypes.byref(nr))
        return [left[i] for i in range(nl.value)], [right[i] for i in range(nr.value)]
