#!/usr/bin/env python3
"""Generate tests/golden/bench_n200.json: the bench workload (BASELINE config 4, clustered synthetic n = 200, diameter
sweep) as committed point sets, plus — for the bounded CPU sample (the D = 2 solve of each sweep) — the cell and state
counts from the instrumented oracle port and the answer of the unmodified reference (oracle/_ref/libadkref.so).

Run in the build container: python tests/golden/make_bench_fixture.py
"""
import importlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.load_package()
synth = importlib.import_module("area_diameter_algorithm_b200.synth")
from oracle.oraclepy import Oracle  # noqa: E402
from oracle.refpy import RefLib, ref_available  # noqa: E402

N, SEED, AMAX, DIAMS = 200, 200, 4.0, [2.0, 3.0, 4.0, 5.0, 6.0]
out = {"generated_by": "tests/golden/make_bench_fixture.py", "n": N, "seed": SEED, "amax": AMAX, "diameters": DIAMS,
       "generator": "area_diameter_algorithm_b200.synth.gaussian_points (python/generate_input_samples.py:37-48 restated)",
       "sets": {}}
O = Oracle()
R = RefLib("x87") if ref_available() else None
for sigma, sample_d in ((3.0, [2.0, 3.0]), (2.0, [2.0, 3.0]), (1.5, [2.0]), (1.0, []), (0.5, [])):
    pts = synth.gaussian_points(N, sigma, SEED)
    x = [p[0] for p in pts]
    y = [p[1] for p in pts]
    entry = {"sigma": sigma, "points": [[repr(p[0]), repr(p[1])] for p in pts], "sample": []}
    for d in sample_d:
        o = O.solve_area_diameter(x, y, None, AMAX, d, True)
        rec = {"dmax": d, "cells": o["cells"], "states": o["states"], "pairs": o["pairs"], "count": o["count"],
               "area": repr(o["area"]), "hull": o["hull"], "diam": o["diam"]}
        if R is not None:
            r = R.solve(RefLib.ALGO_ANTIPODAL_OPTIMIZED, x, y, None, AMAX, d, True)
            assert (r["count"], r["hull"], r["diam"]) == (o["count"], o["hull"], o["diam"]), (r, o)
            rec["ref_entries"] = r["entries_required"]
            rec["ref_solve_ms_build_container"] = r["solve_ms"]
        entry["sample"].append(rec)
        print(sigma, d, rec["cells"], rec["states"], rec.get("ref_solve_ms_build_container"), flush=True)
    out["sets"][str(sigma)] = entry
with open(os.path.join(HERE, "bench_n200.json"), "w") as f:
    json.dump(out, f)
