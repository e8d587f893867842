"""Development probe: per-call latency of small solves (BASELINE configs 1 and 2, n = 50) through the C ABI with host buffers."""
import json, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import __graft_entry__ as ge
pkg = ge.load_package()
s = pkg.Solver(0)
P = json.load(open("tests/golden/points.json"))
for name in ("tests/50/0", "tests/100/0", "uniform/10/0"):
    x = [p[0] for p in P[name]]; y = [p[1] for p in P[name]]
    for what, fn in (("A  A=4", lambda: s.solve_area(x, y, None, 4.0, True)), ("AD A=4 D=4", lambda: s.solve_area_diameter(x, y, None, 4.0, 4.0, True))):
        fn(); fn()
        t0 = time.perf_counter()
        for _ in range(20): r = fn()
        dt = (time.perf_counter() - t0) / 20
        print(f"{name:14s} n={len(x):4d} {what:12s} {dt*1e3:8.3f} ms per call  (kernel_ms {r['kernel_ms']:.3f}, count {r['count']})")
