#!/usr/bin/env python3
"""One-off soak of the oracle (oracle/adk_oracle.c, binary64) against the UNMODIFIED reference (oracle/_ref/libadkref.so, x87) on
seeded random inputs: both solves (with hull, diameter pair and BenchmarkInfo counters), the clockwise orders and the count tables.
Uniform and clustered (Gaussian) point sets, n = 3..60, random K / A / D including binding ones.  BUILD container only.
    python tools/oracle_soak.py --trials 4000 --procs 7 > profiles/r2_oracle_soak.txt"""
import argparse
import os
import random
import sys
import time
from multiprocessing import get_context

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def one(seed):
    from oracle.oraclepy import Oracle
    from oracle.refpy import RefLib
    from util import same_cycle
    o, ref = Oracle(), RefLib("x87")
    rng = random.Random(seed)
    n = rng.choice([3, 4, 5, 8, 12, 20, 30, 45, 60])
    if rng.random() < 0.5:
        x = [rng.uniform(0, 20) for _ in range(n)]; y = [rng.uniform(0, 20) for _ in range(n)]
    else:
        s = rng.choice([0.5, 1.5, 3.0])
        x = [min(20.0, max(0.0, rng.gauss(10, s))) for _ in range(n)]; y = [min(20.0, max(0.0, rng.gauss(10, s))) for _ in range(n)]
    k = rng.choice([None, None, 3, 5, n // 2 + 1, n])
    a = rng.choice([0.5, 2.0, 4.0, 30.0, float("inf")])
    d = rng.choice([1.0, 2.0, 4.0, 8.0, 1e9])
    bad = []

    def same(p, q, what, hull=True):
        ok = p["found"] == q["found"]
        if ok and q["found"]:
            ok = (p["count"] == q["count"] and abs(p["area"] - q["area"]) <= 1e-9 * abs(q["area"]) and sorted(p["diam"]) == sorted(q["diam"])
                  and (not hull or same_cycle(p["hull"], q["hull"])))
        ok = ok and (p["entries_alloc"], p["entries_required"]) == (q["entries_alloc"], q["entries_required"])
        if not ok:
            bad.append((what, seed))
    same(o.solve_area(x, y, k, a, True), ref.solve(0, x, y, k, a, float("inf"), True), "A")
    same(o.solve_area_diameter(x, y, k, a, d, True), ref.solve(2, x, y, k, a, d, True), "AD")
    if o.cw_order(x, y) != ref.cw_order(x, y):
        bad.append(("cw_order", seed))
    if o.count_tables(x, y) != ref.count_tables(x, y):
        bad.append(("count_tables", seed))
    return n, bad


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=4000)
    ap.add_argument("--procs", type=int, default=7)
    ap.add_argument("--seed0", type=int, default=100000)
    a = ap.parse_args()
    t0 = time.time()
    with get_context("spawn").Pool(a.procs) as pool:
        res = pool.map(one, range(a.seed0, a.seed0 + a.trials), chunksize=8)
    bad = [b for _, bs in res for b in bs]
    by_n = {}
    for n, _ in res:
        by_n[n] = by_n.get(n, 0) + 1
    print(f"oracle (binary64) vs unmodified reference (x87): {a.trials} seeded random instances (seeds {a.seed0}..{a.seed0 + a.trials - 1}), "
          f"4 comparisons each (area solve, area+diameter solve incl. hull / diameter pair / BenchmarkInfo counters, clockwise orders, count tables)")
    print("instances by n:", dict(sorted(by_n.items())))
    print(f"differences: {len(bad)}", bad[:20])
    print(f"wall {time.time() - t0:.0f} s on {a.procs} processes")


if __name__ == "__main__":
    main()
