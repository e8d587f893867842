#!/usr/bin/env python3
"""Freeze the UNMODIFIED reference (oracle/_ref/libadkref.so, x87 long double) on the instances bench.py actually times
(run in the BUILD container only; hours of CPU, one process per job).

  bench.sigma1.5.D{3,4,5,6} : the headline sweep beyond the D = 2 sample (tests/golden/bench_n200.json holds D = 2)
  real.{1..5}.D4            : the cell samples at the real call shape (quality_of_results_comparison.py:194-197)
  bench.sigma{1.0,0.5}.D2   : the D = 2 solve of the denser sets of the bench fixture
  bench.sigma1.0.D3         : the solve at which the sigma = 1.0 sweep reaches its final answer
  bench.sigma{2.0,3.0}.D{4,5,6} : the wider sets of the fixture beyond their D = 2, 3 samples
  area.n1000.K16            : BASELINE config 3's point set (synth.gaussian_points(1000, 2.0, 1000)), area path, K = 16

Each job writes tests/golden/heavy/<tag>.json when it ends; `--merge` folds whatever is there into
tests/golden/heavy_outputs.json (the file the GPU tests and bench.py read).
Usage: python tests/golden/make_heavy_golden.py [--jobs N] [--only TAG ...] | --merge
"""
import argparse
import importlib
import json
import os
import sys
import time
from multiprocessing import get_context

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "heavy")
sys.path.insert(0, ROOT)

KINF = 2**64 - 1


def job_list():
    import __graft_entry__ as ge
    ge.load_package()
    synth = importlib.import_module("area_diameter_algorithm_b200.synth")
    with open(os.path.join(HERE, "points.json")) as f:
        allp = json.load(f)
    with open(os.path.join(HERE, "bench_n200.json")) as f:
        fx = json.load(f)
    jobs = []
    for r in (1, 2, 3, 4, 5):
        p = allp[f"real/{r}/0"]
        jobs.append({"tag": f"real.{r}.D4", "name": f"real/{r}/0", "algo": 2, "kmax": KINF, "amax": 4.0, "dmax": 4.0,
                     "x": [q[0] for q in p], "y": [q[1] for q in p]})
    p = synth.gaussian_points(1000, 2.0, 1000)
    jobs.append({"tag": "area.n1000.K16", "name": "synth.gaussian_points(1000,2.0,1000)", "algo": 0, "kmax": 16, "amax": 4.0,
                 "dmax": "inf", "x": [q[0] for q in p], "y": [q[1] for q in p]})
    pts = fx["sets"]["1.5"]["points"]
    for d in (6.0, 5.0, 4.0, 3.0):
        jobs.append({"tag": f"bench.sigma1.5.D{int(d)}", "name": "bench_n200/1.5", "algo": 2, "kmax": KINF, "amax": fx["amax"],
                     "dmax": d, "x": [float(q[0]) for q in pts], "y": [float(q[1]) for q in pts]})
    # the D = 2 solve of the two denser sets SURVEY 8d names (sigma = 1.0, 0.5): the only solves of those sweeps a CPU can afford
    for sigma in ("1.0", "0.5"):
        pts = fx["sets"][sigma]["points"]
        jobs.append({"tag": f"bench.sigma{sigma}.D2", "name": f"bench_n200/{sigma}", "algo": 2, "kmax": KINF, "amax": fx["amax"],
                     "dmax": 2.0, "x": [float(q[0]) for q in pts], "y": [float(q[1]) for q in pts]})
    # sigma = 1.0, D = 3: the first diameter at which the denser sweep reaches its final answer (110 points); ~3.4e9 hash entries,
    # a couple of hours of one core
    pts = fx["sets"]["1.0"]["points"]
    jobs.append({"tag": "bench.sigma1.0.D3", "name": "bench_n200/1.0", "algo": 2, "kmax": KINF, "amax": fx["amax"],
                 "dmax": 3.0, "x": [float(q[0]) for q in pts], "y": [float(q[1]) for q in pts]})
    # the wider sets of the bench fixture (sigma = 2.0, 3.0) beyond the D = 2, 3 samples bench_n200.json holds: minutes each
    for sigma in ("2.0", "3.0"):
        pts = fx["sets"][sigma]["points"]
        for d in (4.0, 5.0, 6.0):
            jobs.append({"tag": f"bench.sigma{sigma}.D{int(d)}", "name": f"bench_n200/{sigma}", "algo": 2, "kmax": KINF, "amax": fx["amax"],
                         "dmax": d, "x": [float(q[0]) for q in pts], "y": [float(q[1]) for q in pts]})
    return jobs


def run(job):
    from oracle.refpy import RefLib
    lib = RefLib("x87")
    t0 = time.time()
    dmax = float("inf") if job["dmax"] == "inf" else job["dmax"]
    r = lib.solve(job["algo"], job["x"], job["y"], job["kmax"], job["amax"], dmax, reconstruct=True)
    rec = {k: job[k] for k in ("tag", "name", "algo", "kmax", "amax", "dmax")}
    rec.update(r)
    rec["n"] = len(job["x"])
    rec["area_hex"] = float(r["area"]).hex() if r["found"] else None
    rec["wall_s"] = time.time() - t0
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, job["tag"] + ".json"), "w") as f:
        json.dump(rec, f)
    return job["tag"], rec["wall_s"]


def merge():
    dest = os.path.join(HERE, "heavy_outputs.json")
    by_tag = {}
    if os.path.exists(dest):                         # runs already merged stay (the per-job files are scratch)
        with open(dest) as f:
            by_tag = {r["tag"]: r for r in json.load(f)["runs"]}
    for fn in sorted(os.listdir(OUT)):
        if fn.endswith(".json"):
            with open(os.path.join(OUT, fn)) as f:
                r = json.load(f)
            by_tag[r["tag"]] = {**by_tag.get(r["tag"], {}), **r}     # keys added later (tests/golden/add_f64_entries.py) stay
    runs = [by_tag[t] for t in sorted(by_tag)]
    with open(dest, "w") as f:
        json.dump({"generated_by": "tests/golden/make_heavy_golden.py",
                   "reference_build": "oracle/_ref/libadkref.so: unmodified reference, g++ -O3, long double = x87 80-bit",
                   "runs": runs}, f, separators=(",", ":"))
    print("merged", len(runs), "runs:", [r["tag"] for r in runs])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, default=5)
    ap.add_argument("--only", nargs="*")
    ap.add_argument("--merge", action="store_true")
    args = ap.parse_args()
    if args.merge:
        return merge()
    jobs = job_list()
    if args.only:
        jobs = [j for j in jobs if j["tag"] in args.only]
    jobs = [j for j in jobs if not os.path.exists(os.path.join(OUT, j["tag"] + ".json"))]
    print("jobs:", [j["tag"] for j in jobs], flush=True)
    with get_context("spawn").Pool(args.jobs) as pool:
        for tag, s in pool.imap_unordered(run, jobs, chunksize=1):
            print(f"done {tag} {s:.0f}s", flush=True)
    merge()


if __name__ == "__main__":
    main()
