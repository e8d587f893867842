#!/usr/bin/env python3
"""Evidence for BASELINE configs 1 and 2 (n = 50, the reference's own CPU-runnable cases) and for the tiled area-only workflow:
per-solve latency through the C ABI with HOST buffers, point sets ROTATING and precompute reuse OFF (every solve recomputes the
count tables and orders, as the reference does), next to the unmodified reference (oracle/_ref) timed on one host core on the
same instances.  Prints one JSON line per config.  usage: python tools/small_configs.py [--reps R]"""
import argparse, importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge

ap = argparse.ArgumentParser(); ap.add_argument("--reps", type=int, default=20); args = ap.parse_args()
pkg = ge.load_package()
synth = importlib.import_module("area_diameter_algorithm_b200.synth")
tiling = importlib.import_module("area_diameter_algorithm_b200.tiling")
from oracle.refpy import RefLib, ref_available
ref = RefLib("x87") if ref_available("x87") else None
s = pkg.Solver(0)
s.set_reuse(False)
sets = [synth.uniform_points(50, 50 + i) for i in range(10)]
pinned = []
for pts in sets:
    hx = torch.tensor([p[0] for p in pts], dtype=torch.float64).pin_memory(); hy = torch.tensor([p[1] for p in pts], dtype=torch.float64).pin_memory()
    pinned.append((hx, hy))

def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps

# ---- config 1: area path, uniform n = 50, A = 4, K = inf, reconstruct (src/bench/main.cpp:57-59,89)
def sweep_a():
    return [s.solve_area_host(hx.data_ptr(), hy.data_ptr(), 50, None, 4.0, True) for hx, hy in pinned]
ra = sweep_a()
t = timed(sweep_a, args.reps) / len(sets)
line = {"config": "1: area-bounded DP, synthetic uniform n=50 (10 point sets rotating, reuse off), A=4, K=inf, reconstruct",
        "ms_per_solve_b200_e2e": 1e3 * t, "solves_per_sec": 1 / t, "nominal_cells_per_solve": 48 * 50 * 49 * 48 // 3,
        "nominal_cells_per_sec": 48 * 50 * 49 * 48 // 3 / t, "answers": [(r["count"], r["area"]) for r in ra]}
if ref:
    t0 = time.perf_counter(); rr = [ref.solve(0, [p[0] for p in pts], [p[1] for p in pts], None, 4.0, float("inf"), True) for pts in sets]
    tr = (time.perf_counter() - t0) / len(sets)
    assert [(r["count"], r["hull"]) for r in rr] == [(r["count"], r["hull"]) for r in ra]
    line.update({"ms_per_solve_reference_1core": 1e3 * tr, "speedup_vs_1core": tr / t})
print(json.dumps(line), flush=True)

# ---- config 2: area+diameter path, same sets, three diameters
for D in (2.0, 4.0, 6.0):
    def sweep_d():
        return [s.solve_area_diameter_host(hx.data_ptr(), hy.data_ptr(), 50, None, 4.0, D, True) for hx, hy in pinned]
    rd = sweep_d()
    t = timed(sweep_d, args.reps) / len(sets)
    line = {"config": f"2: area+diameter-bounded DP, synthetic uniform n=50 (10 point sets rotating, reuse off), A=4, D={D}, K=inf, reconstruct",
            "ms_per_solve_b200_e2e": 1e3 * t, "solves_per_sec": 1 / t, "cells_per_solve": sum(r["cells_executed"] for r in rd) / len(rd),
            "answers": [(r["count"], r["area"]) for r in rd]}
    if ref:
        t0 = time.perf_counter(); rr = [ref.solve(2, [p[0] for p in pts], [p[1] for p in pts], None, 4.0, D, True) for pts in sets]
        tr = (time.perf_counter() - t0) / len(sets)
        assert [(r["found"], r["count"], r["hull"]) for r in rr] == [(r["found"], r["count"], r["hull"]) for r in rd]
        line.update({"ms_per_solve_reference_1core": 1e3 * tr, "speedup_vs_1core": tr / t})
    print(json.dumps(line), flush=True)

# ---- the tiled area-only workflow: windows generated on the device, one C call per slide
with open(os.path.join(ROOT, "tests", "golden", "points.json")) as f:
    allp = json.load(f)
with open(os.path.join(ROOT, "tests", "golden", "tiled_area.json")) as f:
    cases = json.load(f)["cases"]
for c in cases:
    pts = allp[c["name"]]
    def run():
        return s.solve_area_tiled(pts, c["patch_size"], c["step_size"], None, c["max_area"])
    r = run()
    assert (r["count"], r["patches"], r["solved"]) == (c["count"], c["patches"], c["solved"])
    t = timed(run, 5)
    line = {"config": f"tiled area-only workflow, {c['name']} (n={len(pts)}), patch {c['patch_size']} step {c['step_size']}, A={c['max_area']}",
            "windows": c["patches"], "windows_solved": c["solved"], "ms_per_slide_b200": 1e3 * t, "windows_per_sec": c["patches"] / t}
    if ref:
        # the same workflow on the host: windows from the vectorised NumPy generator of tiling.py (the reference's create_patch_dict
        # filters a DataFrame per window and is much slower), the visited windows solved by the unmodified reference on one core
        t0 = time.perf_counter()
        xy, keep, order, member = tiling.make_patches(pts, c["patch_size"], c["step_size"])
        best = None; solved = 0
        for w in order:
            m = member(w)
            if best is not None and best["count"] > len(m): break
            solved += 1
            o = ref.solve(0, xy[m][:, 0].tolist(), xy[m][:, 1].tolist(), None, c["max_area"], float("inf"), True) if len(m) >= 3 else {"found": 0}
            if o["found"] and (best is None or best["count"] < o["count"]): best = o
        tr = time.perf_counter() - t0
        assert solved == c["solved"] and best["count"] == c["count"]
        line.update({"ms_per_slide_host_numpy_plus_reference_1core": 1e3 * tr, "speedup": tr / t})
    print(json.dumps(line), flush=True)
