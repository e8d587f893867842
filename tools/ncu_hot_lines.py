#!/usr/bin/env python3
"""Attribute the per-SASS-instruction metrics of an ncu report to CUDA source lines.
ncu's CSV source page carries metrics only in SASS view; nvdisasm -g gives the line of every SASS instruction of the same
cubin, in the same order.  Usage: ncu_hot_lines.py <report.ncu-rep> <libadk.so> <kernel substring> [top N]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

rep, lib, kern = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H = rows[1]
ii, isamp, isrc = H.index("Instructions Executed"), H.index("# Samples"), H.index("Source")
sass = [(r[isrc].strip(), int(r[ii]), int(r[isamp])) for r in rows[2:] if len(r) >= len(H) and r[ii].isdigit()]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines = None
for f in sorted(os.listdir(tmp)):
    if not f.endswith(".cubin") or "-" in f:
        continue
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    if kern not in txt:
        continue
    cur, infn, out = None, False, []
    for ln in txt.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            infn = kern in m.group(1)
            continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
        if m:
            out.append((m.group(1).strip(), cur))
    if out:
        lines = out
        break
if lines is None:
    sys.exit("kernel not found in cubins")
print(f"# {len(sass)} SASS instructions in the report, {len(lines)} in the cubin")
agg = collections.defaultdict(lambda: [0, 0])
n = min(len(sass), len(lines))
for (s, inst, samp), (_, loc) in zip(sass[:n], lines[:n]):
    agg[loc][0] += inst
    agg[loc][1] += samp
ti = sum(v[0] for v in agg.values()) or 1
ts = sum(v[1] for v in agg.values()) or 1
src_cache = {}
def src(loc):
    if loc is None:
        return ""
    f, l = loc
    for d in ("area-diameter-algorithm_b200/csrc",):
        p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), d, f)
        if os.path.exists(p):
            if p not in src_cache:
                src_cache[p] = open(p).read().splitlines()
            return src_cache[p][l - 1].strip()[:100] if l - 1 < len(src_cache[p]) else ""
    return ""
print(f"# total warp instructions {ti}, stall samples {ts}")
for loc, (inst, samp) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{100 * samp / ts:5.1f}% samples {100 * inst / ti:5.1f}% inst  {loc[0] if loc else '?'}:{loc[1] if loc else 0}  {src(loc)}")

# per-phase totals: a phase starts at a "// ---- " marker or at a function definition in the source file
import bisect
phase_marks = {}
def phase_of(loc):
    if loc is None:
        return "?"
    f, l = loc
    if f not in phase_marks:
        marks = []
        for d in ("area-diameter-algorithm_b200/csrc",):
            pth = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), d, f)
            if os.path.exists(pth):
                for i, ln in enumerate(open(pth).read().splitlines(), 1):
                    t = ln.strip()
                    if t.startswith("// ---- ") or re.match(r"^(template <.*>\s*)?(__device__|__global__|__host__)", t):
                        marks.append((i, t[:70]))
        phase_marks[f] = marks
    marks = phase_marks[f]
    k = bisect.bisect_right([m[0] for m in marks], l) - 1
    return f"{f}:{marks[k][0]} {marks[k][1]}" if k >= 0 else f
pagg = collections.defaultdict(lambda: [0, 0])
for loc, (inst, samp) in agg.items():
    ph = phase_of(loc)
    pagg[ph][0] += inst
    pagg[ph][1] += samp
print("# by phase")
for ph, (inst, samp) in sorted(pagg.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{100 * samp / ts:5.1f}% samples {100 * inst / ti:5.1f}% inst  {ph}")
