#!/usr/bin/env bash
# quick GPU check: parity suite + short bench line (no CPU leg).  usage: r1_quick.sh <tag>
TAG=${1:-q}
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log)
(timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?" >> gpurun_out/${TAG}_bench.err)
[ -x tools/issue_peaks ] && tools/issue_peaks > gpurun_out/issue_peaks.json 2>&1
tail -3 gpurun_out/${TAG}_pytest.log; cut -c1-400 gpurun_out/${TAG}_bench.json; tail -2 gpurun_out/${TAG}_bench.err; cat gpurun_out/issue_peaks.json
