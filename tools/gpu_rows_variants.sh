#!/usr/bin/env bash
# development: config-3 bench for every k_area_rows build variant under build/variants/rows_*.so (+ area parity for two of them)
# usage: tools/gpu_rows_variants.sh <tag>
TAG=${1:-rv}
mkdir -p gpurun_out
one() { # name lib extra-env
  n=$1; so=$2; shift 2
  env "$@" ADK_LIB_PATH=$PWD/$so timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err
  python - "$n" gpurun_out/${TAG}_$n.json <<'PY'
import json,sys,hashlib
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); a=d["answers"][0]
    print(sys.argv[1], "cells/step %.4g" % (d["value"]*d["ms_per_step"]/1e3), "ms/solve %.1f" % d["ms_per_step"], "rows kernel %.1f ms" % d["roofline"]["avg_launch_ms"], "frac %.3f" % d["roofline"]["frac"], "count", a["count"], "area %.12g" % a["area"], hashlib.md5(str(a["hull"]).encode()).hexdigest()[:8])
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
}
for so in build/variants/rows_*.so; do one $(basename $so .so) $so X=1; done
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log)
tail -3 gpurun_out/${TAG}_pytest_gpu.log
