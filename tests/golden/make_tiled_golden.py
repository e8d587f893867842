#!/usr/bin/env python3
"""Freezes outputs of the reference's tiled area-only workflow (external/Area-Selector/run/quality_of_results_comparison.py:77-103:
create_patch_dict + the loop over patches by population with early stop) into tests/golden/tiled_area.json.

Runs in the build container only (needs /root/reference): the patching is the reference's own create_patch_dict
(src/convex_area_selector/convex_polygon_finder.py:139-219, imported as is; matplotlib, which the container lacks and the function
does not use, is stubbed), the solver is the unmodified reference library (oracle/_ref/libadkref.so) in place of the pybind module."""
import importlib
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference/external/Area-Selector"
for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "shapely", "shapely.geometry", "shapely.ops"):
    if name not in sys.modules:
        try:
            importlib.import_module(name)
        except Exception:
            sys.modules[name] = types.ModuleType(name)
sys.path.insert(0, REF)
import pandas as pd                                                     # noqa: E402
from src.convex_area_selector.convex_polygon_finder import create_patch_dict  # noqa: E402
from oracle.refpy import RefLib                                         # noqa: E402


def run_area_only_tiled(lib, points, settings, max_count=None):
    """quality_of_results_comparison.py:77-103, with thesis.AreaOnlyAlgorithm -> the reference library."""
    df = pd.DataFrame()
    df["x_absolute"] = [p[0] for p in points]
    df["y_absolute"] = [p[1] for p in points]
    patch_dict, ordered = create_patch_dict(df, settings)
    result, solved = None, 0
    for key in ordered:
        pp = patch_dict[key]
        if result is not None and result[1] > len(pp):
            break
        r = lib.solve(RefLib.ALGO_EPPSTEIN, [p[0] for p in pp], [p[1] for p in pp], max_count, settings["max_area"], float("inf"), True)
        solved += 1
        if r["found"] and (result is None or result[1] < r["count"]):
            real = []
            for i in r["hull"]:
                for j, p in enumerate(points):
                    if tuple(p) == tuple(pp[i]):
                        real.append(j)
                        break
            result = (r["area"], r["count"], real)
    return result, len(ordered), solved


def main():
    with open(os.path.join(ROOT, "tests", "golden", "points.json")) as f:
        pts = json.load(f)
    lib = RefLib("x87")
    out = {"generated_by": "tests/golden/make_tiled_golden.py", "cases": []}
    for name in ("real/9/0", "real/6/0", "real/3/0"):
        for step in (3 * 0.66, 0.5):
            settings = {"patch_size": 3, "step_size": step, "should_scale": False, "verbose": False, "max_area": 4.0,
                        "microns_per_pixel": 0.23}
            points = [tuple(p) for p in pts[name]]
            res, npatches, solved = run_area_only_tiled(lib, points, settings)
            out["cases"].append({"name": name, "patch_size": 3, "step_size": step, "max_area": 4.0, "patches": npatches, "solved": solved,
                                 "found": res is not None, "area": res[0] if res else None, "count": res[1] if res else None,
                                 "hull": res[2] if res else None})
            print(out["cases"][-1])
    with open(os.path.join(ROOT, "tests", "golden", "tiled_area.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
