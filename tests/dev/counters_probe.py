"""Development probe: BenchmarkInfo counters of the GPU path vs the reference's frozen outputs."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import __graft_entry__ as ge
pkg = ge.load_package()
s = pkg.Solver(0)
g = json.load(open("tests/golden/ref_outputs.json")); P = json.load(open("tests/golden/points.json"))
bad = {}
for r in g["runs"]:
    if len(P[r["name"]]) > 200 or r["algo"] == 1: continue
    x = [p[0] for p in P[r["name"]]]; y = [p[1] for p in P[r["name"]]]
    k = None if r["kmax"] >= 2**63 else r["kmax"]
    a = float(r["amax"]); d = float(r["dmax"])
    o = s.solve_area(x, y, k, a, True, True) if r["algo"] == 0 else s.solve_area_diameter(x, y, k, a, d, True, True)
    key = (r["algo"], r["tag"])
    st = bad.setdefault(key, [0, 0, 0, None])
    st[0] += 1
    st[1] += o["entries_alloc"] != r["entries_alloc"]
    if o["entries_required"] != r["entries_required"]:
        st[2] += 1
        if st[3] is None: st[3] = (r["name"], o["entries_required"], r["entries_required"], o["entries_alloc"], r["entries_alloc"])
for k, v in sorted(bad.items()): print(k, "runs", v[0], "alloc mismatches", v[1], "required mismatches", v[2], v[3])
