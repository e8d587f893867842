"""Shared helpers for the parity tests."""
import json
import math
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REL_TOL = 1e-9   # north_star: hull area within 1e-9 relative


def xy(points):
    return [p[0] for p in points], [p[1] for p in points]


def same_cycle(a, b):
    """Hull index lists equal up to rotation, same direction (src/test/main.cpp:51-66)."""
    if len(a) != len(b):
        return False
    if not a:
        return True
    if b[0] not in a:
        return False
    i = a.index(b[0])
    return all(a[(i + t) % len(a)] == b[t] for t in range(len(b)))


def assert_same_solution(got, want, what="", check_hull=True):
    assert got["found"] == want["found"], (what, got, want)
    if not want["found"]:
        return
    assert got["count"] == want["count"], (what, got["count"], want["count"])
    assert abs(got["area"] - want["area"]) <= REL_TOL * abs(want["area"]), (what, got["area"], want["area"])
    if check_hull:
        assert same_cycle(got["hull"], want["hull"]), (what, got["hull"], want["hull"])
    if want.get("diam") is not None:
        assert got["diam"] is not None and sorted(got["diam"]) == sorted(want["diam"]), (what, got["diam"], want["diam"])


def float32_equal(a, b):
    """EXPECT_FLOAT_EQ: within 4 float32 ULPs."""
    import struct
    fa = struct.unpack("<i", struct.pack("<f", a))[0]
    fb = struct.unpack("<i", struct.pack("<f", b))[0]
    return abs(fa - fb) <= 4


def polygon_area(x, y, hull):
    s = 0.0
    for a, b in zip(hull, hull[1:] + hull[:1]):
        s += x[a] * y[b] - x[b] * y[a]
    return abs(s) / 2


def cross(ax, ay, bx, by, cx, cy):
    return (bx - ax) * (cy - ay) - (cx - ax) * (by - ay)


def check_witness(x, y, sol, amax, dmax=None):
    """Size-independent validity of a returned witness: convex cycle, area, enclosed count, diameter bound."""
    hull = sol["hull"]
    h = len(hull)
    assert h >= 3 and len(set(hull)) == h
    area = polygon_area(x, y, hull)
    assert abs(area - sol["area"]) <= 1e-7 * max(1.0, area), (area, sol["area"])
    assert sol["area"] <= amax * (1 + 1e-12)
    sign = 0
    for t in range(h):
        a, b, c = hull[t], hull[(t + 1) % h], hull[(t + 2) % h]
        cr = cross(x[a], y[a], x[b], y[b], x[c], y[c])
        if abs(cr) > 2e-10:
            assert sign == 0 or (cr > 0) == (sign > 0), "witness is not convex"
            sign = 1 if cr > 0 else -1
    inside = 0
    for p in range(len(x)):
        ok = True
        for t in range(h):
            a, b = hull[t], hull[(t + 1) % h]
            cr = cross(x[a], y[a], x[b], y[b], x[p], y[p])
            if cr * sign < -2e-10:
                ok = False
                break
        inside += ok
    assert inside == sol["count"], (inside, sol["count"])
    if dmax is not None and math.isfinite(dmax):
        far = max(math.dist((x[a], y[a]), (x[b], y[b])) for a in hull for b in hull)
        assert far <= dmax * (1 + 1e-12)


def write_reference_layout(golden, dest):
    """Re-creates data/samples/tests/<n>/{points_i.json,results.json} from the committed golden vectors."""
    for n in range(10, 101, 10):
        d = os.path.join(dest, "tests", str(n))
        os.makedirs(d, exist_ok=True)
        results = []
        for i in range(10):
            name = f"tests/{n}/{i}"
            pts = golden["points"][name]
            with open(os.path.join(d, f"points_{i}.json"), "w") as f:
                json.dump({"points": [{"x": p[0], "y": p[1]} for p in pts], "count": len(pts)}, f)
            t = golden["ref_tests"][name]
            r = {"id": f"Eppstein/Uniform/{n}/iteration:{i}", "max_count": t["max_count"], "max_area": t["max_area"]}
            if t["convex_area"]:
                r["convex_area"] = t["convex_area"]
            results.append(r)
        with open(os.path.join(d, "results.json"), "w") as f:
            json.dump({"count": len(results), "results": results}, f)


def write_experiments_layout(golden, dest, folders):
    """Re-creates data/samples/experiments/<kind>/<idx>/points_{0..99}.json for the reference's bench driver
    (src/bench/main.cpp:28-33,78): the committed golden point sets of that folder, repeated cyclically to 100 files."""
    for kind, idx in folders:
        names = sorted((k for k in golden["points"] if k.startswith(f"{kind}/{idx}/")), key=lambda k: int(k.rsplit("/", 1)[1]))
        assert names, (kind, idx)
        d = os.path.join(dest, "experiments", kind, str(idx))
        os.makedirs(d, exist_ok=True)
        for i in range(100):
            pts = golden["points"][names[i % len(names)]]
            with open(os.path.join(d, f"points_{i}.json"), "w") as f:
                json.dump({"points": [{"x": p[0], "y": p[1]} for p in pts], "count": len(pts)}, f)
    return dest
