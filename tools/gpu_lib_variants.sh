#!/usr/bin/env bash
# development: config-3 bench for every library build under build/variants/*.so   usage: tools/gpu_lib_variants.sh <tag>
TAG=${1:-lv}
mkdir -p gpurun_out
for so in build/variants/*.so; do
  n=$(basename $so .so)
  ADK_LIB_PATH=$PWD/$so timeout 200 python bench.py --workload area_n1000 --steps 4 --warmup 2 --no-cpu > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err
  python - "$n" gpurun_out/${TAG}_$n.json <<'PY'
import json,sys,hashlib
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); a=d["answers"][0]
    print(sys.argv[1], "ms/solve %.1f" % d["ms_per_step"], "rows kernel %.1f ms" % d["roofline"]["avg_launch_ms"], "rest %.1f ms" % (d["ms_per_step"] - d["roofline"]["avg_launch_ms"]), "count", a["count"], "area %.12g" % a["area"], hashlib.md5(str(a["hull"]).encode()).hexdigest()[:8])
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
