"""Development probe: wall time of the partial and the finish of a config-3 solve (area path, n = 1000), separately."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load_package()
synth = importlib.import_module("area_diameter_algorithm_b200.synth")
pts = synth.gaussian_points(1000, 2.0, 1000)
x = [p[0] for p in pts]; y = [p[1] for p in pts]
n = len(x)
s = pkg.Solver(0); s.set_reuse(False)
d_x = torch.tensor(x, dtype=torch.float64, device="cuda"); d_y = torch.tensor(y, dtype=torch.float64, device="cuda")
for it in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    part, st = s.area_partial_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, 4.0, 0, 1)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    r = s.area_finish_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, 4.0, part, True)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(it, "partial %.1f ms (device %.1f, rows kernel %.1f, launches %d)  finish %.2f ms  count %d hull %s" %
          (1e3 * (t1 - t0), st["kernel_ms"], st["dp_ms"], st["launches"], 1e3 * (t2 - t1), r["count"], r["hull"][:4]), flush=True)
