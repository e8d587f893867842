"""Development probe: device time of the angular-order stage (k_angles + k_order_sort, both modes) at n = 1000 .. 4096."""
import os, sys, time, random
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
pkg = ge.load_package()
s = pkg.Solver(0)
s.set_reuse(False)
for n in (1000, 2048, 4096):
    rng = random.Random(n)
    x = [rng.uniform(0, 20) for _ in range(n)]; y = [rng.uniform(0, 20) for _ in range(n)]
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        s.cw_order(x, y)                      # includes the n*(n-1) D2H copy of the lists
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("n", n, "cw_order call (kernels + %d MB D2H): %.2f ms" % (4 * n * (n - 1) // 2**20, 1e3 * dt), flush=True)
