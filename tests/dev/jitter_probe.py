"""Development probe: per-solve wall / device / DP-kernel times of the headline sweep, to see where run-to-run jitter comes from."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load_package()
fx = json.load(open(os.path.join(ROOT, "tests", "golden", "bench_n200.json")))
pts = fx["sets"]["1.5"]["points"]
x = [float(p[0]) for p in pts]; y = [float(p[1]) for p in pts]
n = len(x)
s = pkg.Solver(0)
s.set_reuse(False)
hx = torch.tensor(x, dtype=torch.float64).pin_memory(); hy = torch.tensor(y, dtype=torch.float64).pin_memory()
for it in range(8):
    row = []
    t_sweep = time.perf_counter()
    for d in (2.0, 3.0, 4.0, 5.0, 6.0):
        t0 = time.perf_counter()
        r = s.solve_area_diameter_host(hx.data_ptr(), hy.data_ptr(), n, None, 4.0, d, True)
        w = 1e3 * (time.perf_counter() - t0)
        row.append("%.1f/%.1f/%.1f" % (w, r["kernel_ms"], r["dp_ms"]))
    print(it, "sweep %.1f ms | wall/dev/dp per solve:" % (1e3 * (time.perf_counter() - t_sweep)), " ".join(row), flush=True)
