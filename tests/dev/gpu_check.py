"""Development check (run under gpurun): CUDA path vs oracle / reference on golden fixtures."""
import json, math, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle.oraclepy import Oracle

pts = json.load(open(os.path.join(ROOT, "tests/golden/points.json")))
reft = json.load(open(os.path.join(ROOT, "tests/golden/ref_tests.json")))
refo = json.load(open(os.path.join(ROOT, "tests/golden/ref_outputs.json")))
S = pkg.Solver(0)
O = Oracle()

def xy(name):
    p = pts[name]
    return [q[0] for q in p], [q[1] for q in p]

def same_cycle(a, b):
    if len(a) != len(b): return False
    if not a: return True
    if b[0] not in a: return False
    i = a.index(b[0])
    return all(a[(i + t) % len(a)] == b[t] for t in range(len(b)))

bad = 0
# stage checks
for name in ["tests/10/0", "tests/30/3", "tests/50/0", "uniform/0/0", "real/9/0", "real/6/0"]:
    x, y = xy(name)
    t0 = time.time(); g = S.count_tables(x, y); t1 = time.time(); o = O.count_tables(x, y)
    ok = g == o
    print("tables", name, "OK" if ok else "MISMATCH", f"{t1-t0:.3f}s", flush=True)
    bad += not ok
    if not ok:
        n = len(x)
        for k in range(3):
            d = [i for i in range(len(g[k])) if g[k][i] != o[k][i]]
            print("  table", k, "diffs", len(d), d[:5], [(g[k][i], o[k][i]) for i in d[:5]])
    g = S.cw_order(x, y); o = O.cw_order(x, y)
    ok = g == o
    print("cw", name, "OK" if ok else "MISMATCH", flush=True)
    bad += not ok
    n = len(x)
    import random
    rng = random.Random(1)
    nb = 0
    for _ in range(60):
        i, j = rng.sample(range(n), 2)
        if y[j] < y[i]: i, j = j, i
        if y[j] < y[i] or i == j: continue
        g = S.slope_order(x, y, i, j); o = O.slope_order(x, y, i, j)
        if g != o:
            nb += 1
            if nb < 3: print("  slope mismatch", i, j, g[:12], o[:12])
    print("slope", name, "OK" if nb == 0 else f"MISMATCH {nb}", flush=True)
    bad += nb > 0
    nb = 0
    for _ in range(40):
        s, t = sorted(rng.sample(range(n), 2))
        g = S.lune_lists(x, y, s, t); o = O.lune_lists(x, y, s, t)
        if list(g[0]) != list(o[0]) or list(g[1]) != list(o[1]):
            nb += 1
            if nb < 3: print("  lune mismatch", s, t, g, o)
    print("lune", name, "OK" if nb == 0 else f"MISMATCH {nb}", flush=True)
    bad += nb > 0

runs = {}
for r in refo["runs"]:
    runs.setdefault((r["name"], r["tag"]), []).append(r)

def check(tag, g, r, name):
    global bad
    ok = g["found"] == r["found"]
    if ok and r["found"]:
        ok = g["count"] == r["count"] and same_cycle(g["hull"], r["hull"]) and abs(g["area"] - r["area"]) <= 1e-9 * abs(r["area"])
        if r["diam"] is not None:
            ok = ok and sorted(g["diam"]) == sorted(r["diam"])
    if not ok:
        bad += 1
        print("MISMATCH", tag, name, {k: g[k] for k in ("found", "count", "area", "hull", "diam")}, {k: r[k] for k in ("found", "count", "area", "hull", "diam")}, flush=True)
    return ok

tA = tAD = 0.0
for n in range(10, 101, 10):
    okc = 0; tot = 0
    for i in range(10):
        name = f"tests/{n}/{i}"
        x, y = xy(name)
        t = reft[name]
        rA = runs[(name, "test.A")][0]
        t0 = time.time()
        g = S.solve_area(x, y, t["max_count"], t["max_area"], True)
        tA += time.time() - t0
        tot += 1; okc += check("A", g, rA, name)
        for tag in ("test.ADopt", "bind.ADopt.50", "bind.ADopt.75"):
            for r in runs.get((name, tag), []):
                dm = float("inf") if r["dmax"] == "inf" else r["dmax"]
                am = float("inf") if r["amax"] == "inf" else r["amax"]
                t0 = time.time()
                g = S.solve_area_diameter(x, y, r["kmax"], am, dm, True)
                tAD += time.time() - t0
                tot += 1; okc += check(tag, g, r, name)
    print(f"fixtures n={n}: {okc}/{tot} ok  (gpu wall A {tA:.2f}s AD {tAD:.2f}s)", flush=True)

for r in refo["runs"]:
    if r["tag"].startswith(("bench.", "real.")):
        name = r["name"]; x, y = xy(name)
        dm = float("inf") if r["dmax"] == "inf" else r["dmax"]
        am = float("inf") if r["amax"] == "inf" else r["amax"]
        t0 = time.time()
        if r["algo"] == 0: g = S.solve_area(x, y, r["kmax"], am, True)
        else: g = S.solve_area_diameter(x, y, r["kmax"], am, dm, True)
        dt = time.time() - t0
        ok = check(r["tag"], g, r, name)
        print(r["tag"], name, "D=", r["dmax"], "ok" if ok else "BAD", f"gpu {dt*1e3:.1f} ms (kernel {g['kernel_ms']:.2f} dp {g['dp_ms']:.2f}) ref {r['solve_ms']:.1f} ms cells {g['cells_executed']} states {g['states']} ref_entries {r['entries_required']}", flush=True)
print("TOTAL BAD", bad)
