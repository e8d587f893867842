#!/usr/bin/env python3
"""Development tool: static SASS instruction count per source line of one kernel (nvdisasm --print-line-info of the cubin).
usage: tools/sass_by_line.py <lib.so> <kernel-substring> [top]"""
import collections, os, re, subprocess, sys, tempfile
lib, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
cnt = collections.Counter(); total = 0
for cub in os.listdir(d):
    out = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cub)], capture_output=True, text=True).stdout
    on = False; cur = None
    for ln in out.splitlines():
        if ln.startswith(".text."):
            on = pat in ln; cur = None; continue
        if ln.startswith(".section") or ln.startswith(".text") : on = on and False
        if not on: continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            cnt[cur] += 1; total += 1
print("total SASS instructions:", total)
src = {}
for (f, l), c in cnt.most_common(top) if cnt else []:
    if f not in src:
        for root in ("area-diameter-algorithm_b200/csrc",):
            p = os.path.join(root, f)
            if os.path.exists(p): src[f] = open(p).read().splitlines()
    text = src.get(f, [""] * (l + 1))[l - 1].strip()[:110] if f in src and l <= len(src[f]) else ""
    print(f"{c:6d}  {f}:{l}  {text}")
