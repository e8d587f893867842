#!/usr/bin/env python3
"""Every solve the bench lines under profiles/ print an answer for, beside the CPU answer that pins it: the unmodified x87 reference
(tests/golden/bench_n200.json, heavy_outputs.json) or, where no CPU can afford the reference, the binary64 oracle with the work of
one solve sharded over processes (oracle_sharded.json, oracle_sharded_area.json).  CPU only; reads committed files.
    python tools/pins_vs_recorded.py > profiles/r2_pins_vs_recorded_answers.txt"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


def last_line(path):
    with open(path) as f:
        return [json.loads(l) for l in f if l.lstrip().startswith("{")][-1]


def main():
    fx = json.load(open(os.path.join(G, "bench_n200.json")))
    heavy = {r["tag"]: r for r in json.load(open(os.path.join(G, "heavy_outputs.json")))["runs"]}
    osh = {r["tag"]: r for r in json.load(open(os.path.join(G, "oracle_sharded.json")))["runs"]}
    rows = []
    for sigma, prof in (("1.5", "r2_bench_n1.json"), ("1.0", "r2_bench_sigma1.0_n1.json"), ("0.5", "r2_bench_sigma0.5_n1.json")):
        line = last_line(os.path.join(ROOT, "profiles", prof))
        for a in line["answers"]:
            d = a["D"]
            tag = f"bench.sigma{sigma}.D{d:g}"
            pins = []
            for smp in fx["sets"][sigma]["sample"]:
                if float(smp["dmax"]) == d:
                    pins.append(("x87 reference", smp["count"], float(smp["area"]), smp["diam"]))
            if tag in heavy:
                pins.append(("x87 reference", heavy[tag]["count"], heavy[tag]["area"], heavy[tag]["diam"]))
            if tag in osh:
                pins.append(("binary64 oracle, sharded", osh[tag]["count"], osh[tag]["area"], osh[tag]["diam"]))
            for who, cnt, area, diam in pins or [("- none -", None, None, None)]:
                ok = cnt is not None and cnt == a["count"] and sorted(diam) == sorted(a["diam"]) and abs(area - a["area"]) <= 1e-9 * area
                rows.append((f"AD n=200 sigma={sigma} D={d:g}", prof, a["count"], repr(a["area"]), tuple(a["diam"]), who,
                             "" if cnt is None else cnt, "" if area is None else repr(area), "EQUAL" if ok else ("-" if cnt is None else "DIFFERENT")))
        held = [osh.get(f"bench.sigma{sigma}.D{d:g}") for d in line["config"]["diameters"]]
        if all(held):
            c, s = sum(h["cells"] for h in held), sum(h["states"] for h in held)
            rows.append((f"AD n=200 sigma={sigma} sweep totals", prof, f"cells {line['config']['cells_per_step']}", f"states {line['config']['states_per_step']}", "",
                         "binary64 oracle, sharded", f"cells {c}", f"states {s}",
                         "EQUAL" if (c, s) == (line["config"]["cells_per_step"], line["config"]["states_per_step"]) else "DIFFERENT"))
    apath = os.path.join(G, "oracle_sharded_area.json")
    if os.path.exists(apath):
        o = json.load(open(apath))["runs"][0]
        a = last_line(os.path.join(ROOT, "profiles", "r2_bench_area_n1000_n1.json"))["answers"][0]
        ok = a["count"] == o["count"] and a["hull"] == o["hull"] and abs(a["area"] - o["area"]) <= 1e-9 * o["area"]
        rows.append(("A n=1000 sigma=2 K=inf (config 3)", "r2_bench_area_n1000_n1.json", a["count"], repr(a["area"]), f"hull of {len(a['hull'])}",
                     f"binary64 oracle, anchors sharded (K={o['kmax']})", o["count"], repr(o["area"]), "EQUAL (hull cycle too)" if ok else "DIFFERENT"))
    cpath = os.path.join(ROOT, "profiles", "r2_bench_cells_batch_n1.json")
    if os.path.exists(cpath):
        ro = json.load(open(os.path.join(G, "ref_outputs.json")))
        pins = {}
        for r in ro["runs"] + ro.get("survey_recorded", []):
            if r.get("algo") == 2 and str(r["name"]).startswith("real/") and float(r["dmax"]) == 4.0 and float(r["amax"]) == 4.0:
                pins[r["name"]] = r
        for i in range(1, 6):
            if f"real.{i}.D4" in heavy:
                pins[f"real/{i}/0"] = heavy[f"real.{i}.D4"]
        for i, a in enumerate(last_line(cpath)["answers"]):
            o = pins.get(f"real/{i}/0")
            ok = o is not None and a["count"] == o["count"] and sorted(a["diam"]) == sorted(o["diam"]) and abs(a["area"] - o["area"]) <= 1e-9 * o["area"]
            rows.append((f"AD cell sample real/{i} n={a['n']} A=4 D=4 (config 5)", "r2_bench_cells_batch_n1.json", a["count"], repr(a["area"]), tuple(a["diam"]),
                         "x87 reference" if o else "- none -", o["count"] if o else "", repr(o["area"]) if o else "", "EQUAL" if ok else ("-" if o is None else "DIFFERENT")))
    print("solve | bench line under profiles/ | B200: count | area | diameter pair | pinned by | count | area | verdict")
    for r in rows:
        print(" | ".join(str(v) for v in r))


if __name__ == "__main__":
    main()
