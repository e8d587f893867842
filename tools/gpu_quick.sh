#!/usr/bin/env bash
# quick GPU check: parity suite + short bench line (no CPU leg).  usage: tools/gpu_quick.sh <tag> [pytest -k filter]
TAG=${1:-q}
FILTER=${2:-}
mkdir -p gpurun_out
if [ -n "$FILTER" ]; then
  (timeout 1200 python -m pytest tests -m gpu -q -x -k "$FILTER" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log)
else
  (timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log)
fi
(timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?" >> gpurun_out/${TAG}_bench.err)
tail -15 gpurun_out/${TAG}_pytest.log; cut -c1-600 gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
