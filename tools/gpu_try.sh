#!/usr/bin/env bash
# development: parity + headline bench of one library variant against the in-tree build.  usage: tools/gpu_try.sh <tag> <variant.so>
TAG=${1:-t}; SO=$2
mkdir -p gpurun_out
(ADK_LIB_PATH=$PWD/$SO timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log); tail -3 gpurun_out/${TAG}_pytest.log
for v in variant default variant default; do
  if [ $v = variant ]; then export ADK_LIB_PATH=$PWD/$SO; else unset ADK_LIB_PATH; fi
  timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/${TAG}_$v.json 2> gpurun_out/${TAG}_$v.err
  python - $v gpurun_out/${TAG}_$v.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); print(sys.argv[1], "%.4g cells/s  %.1f ms/step  kernel %.2f ms/launch" % (d["value"], d["ms_per_step"], d["roofline"]["avg_launch_ms"]), d["answers_checked_against_frozen_reference"])
except Exception as e: print(sys.argv[1], "FAILED", e)
PY
done
