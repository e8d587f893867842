#!/usr/bin/env bash
# development: short bench of every library variant under build/variants/   usage: tools/gpu_variants.sh <tag>
TAG=${1:-v}
mkdir -p gpurun_out
for so in build/variants/*.so; do
  n=$(basename $so .so)
  ADK_LIB_PATH=$PWD/$so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err
  python - "$n" gpurun_out/${TAG}_$n.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(sys.argv[1], "%.4g cells/s  %.1f ms/step" % (d["value"], d["ms_per_step"]))
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
