import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
from importlib import import_module
synth = import_module("area_diameter_algorithm_b200.synth")
S = pkg.Solver(0)
for std in (3.0, 2.0, 1.5, 1.0, 0.5):
    pts = synth.gaussian_points(200, std, 200)
    x = [p[0] for p in pts]; y = [p[1] for p in pts]
    for D in (2.0, 3.0, 4.0, 5.0, 6.0):
        t0 = time.time()
        try:
            g = S.solve_area_diameter(x, y, None, 4.0, D, True)
        except Exception as e:
            print("std", std, "D", D, "ERR", e, flush=True); continue
        dt = time.time() - t0
        print(f"std {std} D {D}: count {g['count']} area {g['area']:.6f} wall {dt*1e3:.1f} ms kernel {g['kernel_ms']:.2f} dp {g['dp_ms']:.2f} cells {g['cells_executed']:.3e} states {g['states']:.3e} pairs {g['pairs']} -> {g['cells_executed']/max(g['dp_ms'],1e-9)/1e6:.2f} Gcell/s", flush=True)
    t0 = time.time(); g = S.solve_area(x, y, None, 4.0, True); dt = time.time() - t0
    print(f"std {std} A: count {g['count']} area {g['area']:.6f} wall {dt*1e3:.1f} ms kernel {g['kernel_ms']:.2f} dp {g['dp_ms']:.2f} nominal {g['cells_nominal']:.3e} executed {g['cells_executed']:.3e}", flush=True)
