"""Development probe for compute-sanitizer: a few small solves of both paths (kept tiny: the tools slow kernels ~100x)."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import __graft_entry__ as ge
pkg = ge.load_package()
s = pkg.Solver(0)
P = json.load(open("tests/golden/points.json"))
for name, a, d in (("tests/30/0", 60.0, 9.0), ("tests/60/1", 80.0, 14.0), ("gaussian/12/0", 4.0, 4.0)):
    x = [p[0] for p in P[name]]; y = [p[1] for p in P[name]]
    r1 = s.solve_area(x, y, None, a, True, True)
    r2 = s.solve_area_diameter(x, y, None, a, d, True, True)
    parts = [s.area_diameter_partial(x, y, None, a, d, k, 3)[0] for k in range(3)]
    r3 = s.area_diameter_finish(x, y, None, a, d, s.merge(parts), True)
    assert (r2["count"], r2["hull"]) == (r3["count"], r3["hull"])
    print(name, r1["count"], r2["count"], r2["cells_executed"])
s.close()
print("done")
