"""Development probe: GPU vs oracle on lattice subsets (many exactly collinear triples, equal coordinates)."""
import sys, os, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import __graft_entry__ as ge
from oracle.oraclepy import Oracle
from util import same_cycle
pkg = ge.load_package(); s = pkg.Solver(0); o = Oracle()
rng = random.Random(99)
bad = 0
for trial in range(60):
    g = rng.choice([4, 5, 6, 7]); n = rng.choice([6, 10, 16, 24, min(36, g * g)])
    n = min(n, g * g)
    cells = rng.sample([(i, j) for i in range(g) for j in range(g)], n)
    sc = rng.choice([1.0, 0.5, 1.7])
    x = [c[0] * sc for c in cells]; y = [c[1] * sc for c in cells]
    k = rng.choice([None, 4, 6]); a = rng.choice([2.0, 6.0, float("inf")]); d = rng.choice([2.0, 3.5, float("inf")])
    for what, fg, fo in (("A", lambda: s.solve_area(x, y, k, a, True), lambda: o.solve_area(x, y, k, a, True)),
                         ("AD", lambda: s.solve_area_diameter(x, y, k, a, d, True), lambda: o.solve_area_diameter(x, y, k, a, d, True))):
        try:
            G = fg()
        except Exception as e:
            print(trial, what, 'GPU raised', e, n, k, a, d, list(zip(x, y))); bad += 1; continue
        try:
            O = fo()
        except Exception as e:
            print(trial, what, 'ORACLE raised', repr(e), n, k, a, d, G['count'], G['hull'], list(zip(x, y))); bad += 1; continue
        ok = G["found"] == O["found"] and (not O["found"] or (G["count"] == O["count"] and abs(G["area"] - O["area"]) <= 1e-9 * max(O["area"], 1e-300)))
        okh = ok and (not O["found"] or (same_cycle(G["hull"], O["hull"]) and (O["diam"] is None or sorted(G["diam"]) == sorted(O["diam"]))))
        if not okh:
            bad += 1
            print(trial, what, "n", n, "k", k, "a", a, "d", d, "count/area ok" if ok else "MISMATCH", (G["count"], G["area"], G["hull"], G["diam"]), (O["count"], O["area"], O["hull"], O["diam"]))
print("mismatches", bad, "of 120")
