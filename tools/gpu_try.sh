#!/usr/bin/env bash
# development: short bench under a few environment settings   usage: tools/gpu_try.sh <tag> "ENV=.. ENV=.." ...
TAG=$1; shift
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_$i.json 2> gpurun_out/${TAG}_$i.err
  python - "$envs" gpurun_out/${TAG}_$i.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(sys.argv[1], "%.4g cells/s  %.1f ms/step" % (d["value"], d["ms_per_step"]))
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
