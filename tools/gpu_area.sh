#!/usr/bin/env bash
# development: area-path parity + config-3 bench under several cluster sizes.  usage: tools/gpu_area.sh <tag>
TAG=${1:-a}
mkdir -p gpurun_out
for CS in 0 2; do
  (ADK_AREA_CLUSTER=$CS timeout 900 python -m pytest tests -m gpu -q -x -k "area_path or config3 or edge_cases or oracle_differential or shards_merge or batch_area or tiled or slope" > gpurun_out/${TAG}_pytest_cs$CS.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_cs$CS.log)
  tail -2 gpurun_out/${TAG}_pytest_cs$CS.log
done
for CS in 0; do
  ADK_AREA_CLUSTER=$CS timeout 120 python bench.py --workload area_n1000 --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_area_cs$CS.json 2> gpurun_out/${TAG}_area_cs$CS.err
  python - $CS gpurun_out/${TAG}_area_cs$CS.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print("CS", sys.argv[1], "ms/solve %.1f" % d["ms_per_step"], "rows kernel %.1f ms" % d["roofline"]["avg_launch_ms"], "frac %.3f" % d["roofline"]["frac"], d["answers"][0]["count"])
except Exception as e: print("CS", sys.argv[1], "FAILED", e)
PY
done
