"""The tiled area-only workflow of the reference's Python driver, on top of the C ABI.

Reference: external/Area-Selector/run/quality_of_results_comparison.py:77-103 (run_area_only_tiled) and
src/convex_area_selector/convex_polygon_finder.py:139-219 (create_patch_dict): the slide is cut into square windows
(patch_size, step_size), the windows are visited by population, largest first, each one is solved with the area-only
algorithm, and the walk stops at the first window that holds fewer points than the best polygon found so far.

Here the windows are generated in one vectorised pass (two boolean membership matrices and one integer product give
every window's population; no DataFrame filtering per window) and solved in waves through adk_solve_batch, so the
walk costs one C call per wave instead of one pybind call plus a list-of-tuples marshalling per window.  The answers
are the reference's: same window grid (np.arange from min - 0.01), same inclusive bounds, same duplicate removal,
same visiting order (population descending, ties in x-major grid order), same early stop and strict improvement.
"""
import numpy as np

from .capi import ALGO_AREA


def make_patches(points, patch_size, step_size):
    """Returns (xy, keep, order, member): the de-duplicated coordinates (first occurrence kept) with their indices in
    `points`, the non-empty windows as (ix, iy) in visiting order, and a function window -> indices into xy."""
    a = np.asarray(points, dtype=np.float64).reshape(-1, 2)
    _, first = np.unique(a, axis=0, return_index=True)          # DataFrame.drop_duplicates(): keep the first of equal rows
    keep = np.sort(first)
    xy = a[keep]
    if len(xy) == 0:
        return xy, keep, [], lambda w: np.zeros(0, dtype=np.int64)
    xmin, xmax, ymin, ymax = xy[:, 0].min(), xy[:, 0].max(), xy[:, 1].min(), xy[:, 1].max()
    xr = np.array([xmin - 0.01]) if (xmax - xmin) < patch_size else np.arange(xmin - 0.01, xmax, step_size)
    yr = np.array([ymin - 0.01]) if (ymax - ymin) < patch_size else np.arange(ymin - 0.01, ymax, step_size)
    mx = (xy[:, 0][:, None] >= xr[None, :]) & (xy[:, 0][:, None] <= (xr + patch_size)[None, :])      # Series.between: inclusive
    my = (xy[:, 1][:, None] >= yr[None, :]) & (xy[:, 1][:, None] <= (yr + patch_size)[None, :])
    pop = mx.T.astype(np.int32) @ my.astype(np.int32)                                                # (nx, ny) populations
    flat = pop.reshape(-1)                                                                           # x-major = itertools.product order
    idx = np.nonzero(flat)[0]
    idx = idx[np.argsort(-flat[idx], kind="stable")]                                                 # sorted(..., reverse=True) is stable
    order = [(int(t // len(yr)), int(t % len(yr))) for t in idx]
    return xy, keep, order, (lambda w: np.nonzero(mx[:, w[0]] & my[:, w[1]])[0])


def solve_area_tiled(solver, points, patch_size=3.0, step_size=3 * 0.66, kmax=None, amax=4.0, wave=8):
    """run_area_only_tiled on the GPU.  Returns a dict {found, area, count, hull (indices into `points`), patches, solved}."""
    xy, keep, order, member = make_patches(points, patch_size, step_size)
    best, solved, pos = None, 0, 0
    stop = False
    while pos < len(order) and not stop:
        # windows of this wave: stop collecting at the first one the early-stop rule already rejects
        wins = []
        for w in order[pos:pos + wave]:
            m = member(w)
            if best is not None and best["count"] > len(m):
                stop = True
                break
            wins.append(m)
        if not wins:
            break
        res = solver.solve_batch([xy[m] for m in wins], ALGO_AREA, kmax, amax, float("inf"), True)
        for m, r in zip(wins, res):
            if best is not None and best["count"] > len(m):      # the rule, re-applied in visiting order inside the wave
                stop = True
                break
            solved += 1
            if r["found"] and (best is None or best["count"] < r["count"]):
                best = {"found": 1, "area": r["area"], "count": r["count"], "hull": [int(keep[m[i]]) for i in r["hull"]]}
        pos += len(wins)
    out = best if best is not None else {"found": 0, "area": 0.0, "count": 0, "hull": []}
    out = dict(out)
    out["patches"] = len(order)
    out["solved"] = solved
    return out
