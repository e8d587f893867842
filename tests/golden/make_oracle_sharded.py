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
DEST_AREA = os.path.join(HERE, "oracle_sharded_area.json")      # a file of its own: bench.py and the AD tests read DEST record by record


def work(job):
    from oracle.oraclepy import Oracle
    x, y, amax, dmax, shard, nshards = job
    return Oracle().solve_area_diameter_shard(x, y, None, amax, dmax, shard, nshards)


def area_work(job):
    from oracle.oraclepy import Oracle
    x, y, kmax, amax, shard, nshards = job
    return Oracle().solve_area_shard(x, y, kmax, amax, True, shard, nshards)


def save(rec, dest=DEST):
    runs = []
    if os.path.exists(dest):
        with open(dest) as f:
            runs = [r for r in json.load(f)["runs"] if r["tag"] != rec["tag"]]
    runs.append(rec)
    with open(dest, "w") as f:
        json.dump({"generated_by": "tests/golden/make_oracle_sharded.py", "runs": sorted(runs, key=lambda r: r["tag"])}, f, indent=1)
    print(rec)


def area_main(a):
    """BASELINE config 3 (bench.py --workload area_n1000): Gaussian n = 1000, sigma = 2, seed 1000, A = 4, area path.  The anchors
    are dealt over processes; K is capped at --kmax (default 180) because the table of the restatement holds K + 1 levels (2.9 GB per
    process at 180): the answer has 177 points, so levels 178..180 come out empty and the capped solve is the K = inf solve."""
    import importlib.util
    from oracle.oraclepy import Oracle
    spec = importlib.util.spec_from_file_location("adk_synth", os.path.join(ROOT, "area-diameter-algorithm_b200", "synth.py"))
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)
    pts = synth.gaussian_points(1000, 2.0, 1000)
    x = [p[0] for p in pts]; y = [p[1] for p in pts]
    t0 = time.time()
    with get_context("spawn").Pool(a.procs) as pool:
        parts = pool.map(area_work, [(x, y, a.kmax, 4.0, s, a.shards) for s in range(a.shards)], chunksize=1)
    m = Oracle.merge_area_shards(parts)
    save({"tag": f"area.n1000.K{a.kmax}", "source": "oracle/adk_oracle.c (binary64 restatement), anchors sharded", "path": "area",
          "points": "synth.gaussian_points(1000, 2.0, 1000)", "n": 1000, "kmax": a.kmax, "amax": 4.0, "found": m["found"], "count": m["count"],
          "area": m["area"], "area_hex": float(m["area"]).hex(), "hull": m["hull"], "diam": m["diam"], "cells": m["cells"],
          "entries_required": m["entries_required"], "shards": a.shards, "wall_s": time.time() - t0,
          "cpu_s": sum(p["solve_ms"] for p in parts) / 1e3}, DEST_AREA)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--area", action="store_true", help="config 3's n = 1000 set on the area path, anchors sharded")
    ap.add_argument("--kmax", type=int, default=180)
    ap.add_argument("--sigma")
    ap.add_argument("--dmax", type=float)
    ap.add_argument("--procs", type=int, default=7)
    ap.add_argument("--shards", type=int, default=56)
    a = ap.parse_args()
    if a.area:
        return area_main(a)
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
    save(rec)


if __name__ == "__main__":
    main()
