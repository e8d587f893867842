#!/usr/bin/env python3
"""Answers of the binary64 oracle (oracle/adk_oracle.c, pinned to the reference in tests/test_oracle_cpu.py) for bench instances
one core cannot afford: the candidate pairs of a solve are dealt over processes (adko_solve_area_diameter_shard) and the shards'
bests merged.  Run in the BUILD container only: python tests/golden/make_oracle_sharded.py --sigma 0.5 --dmax 2 [--procs 7]
Appends to tests/golden/oracle_sharded.json (read by bench.py for the diameters that have no frozen reference output: count,
diameter pair and area are compared; these are ORACLE answers, labelled as such, not reference outputs)."""
import argparse
import json
import os
import sys
import time
from multiprocessing import get_context

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
DEST = os.path.join(HERE, "oracle_sharded.json")


def work(job):
    from oracle.oraclepy import Oracle
    x, y, amax, dmax, shard, nshards = job
    return Oracle().solve_area_diameter_shard(x, y, None, amax, dmax, shard, nshards)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sigma", required=True)
    ap.add_argument("--dmax", type=float, required=True)
    ap.add_argument("--procs", type=int, default=7)
    ap.add_argument("--shards", type=int, default=56)
    a = ap.parse_args()
    from oracle.oraclepy import Oracle
    with open(os.path.join(HERE, "bench_n200.json")) as f:
        fx = json.load(f)
    pts = fx["sets"][str(float(a.sigma))]["points"]
    x = [float(p[0]) for p in pts]; y = [float(p[1]) for p in pts]
    t0 = time.time()
    with get_context("spawn").Pool(a.procs) as pool:
        parts = pool.map(work, [(x, y, fx["amax"], a.dmax, s, a.shards) for s in range(a.shards)], chunksize=1)
    m = Oracle.merge_shards(parts)
    rec = {"tag": f"bench.sigma{float(a.sigma)}.D{a.dmax:g}", "source": "oracle/adk_oracle.c (binary64 restatement), candidate pairs sharded",
           "sigma": float(a.sigma), "dmax": a.dmax, "amax": fx["amax"], "found": m["found"], "count": m["count"], "area": m["area"],
           "area_hex": float(m["area"]).hex(), "diam": m["diam"], "cells": m["cells"], "states": m["states"], "pairs": m["pairs"],
           "shards": a.shards, "wall_s": time.time() - t0, "cpu_s": sum(p["solve_ms"] for p in parts) / 1e3}
    runs = []
    if os.path.exists(DEST):
        with open(DEST) as f:
            runs = [r for r in json.load(f)["runs"] if r["tag"] != rec["tag"]]
    runs.append(rec)
    with open(DEST, "w") as f:
        json.dump({"generated_by": "tests/golden/make_oracle_sharded.py", "runs": sorted(runs, key=lambda r: r["tag"])}, f, indent=1)
    print(rec)


if __name__ == "__main__":
    main()
