#!/usr/bin/env bash
# headline bench at N = 8, 4, 2 ranks (torchrun) on one 8-GPU box + the multi-GPU parity tests.  usage: tools/gpu_scale.sh <tag>
TAG=${1:-r2}
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -q -x -k "multi_gpu or several_gpus" > gpurun_out/${TAG}_pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_multi.log)
tail -3 gpurun_out/${TAG}_pytest_multi.log
for N in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N bench.py --gpus $N --steps 5 --warmup 3 \
     > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
  echo "N=$N exit $?"
  python - gpurun_out/${TAG}_bench_n$N.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["n_gpus"], "value %.4g" % d["value"], "ms/step %.1f" % d["ms_per_step"], "e2e(multi) %.4g %.1f ms" % (d["e2e"]["value"], d["e2e"]["ms_per_step"]), "share", round(d["roofline"]["share_of_step"],3), "exch", d["e2e"].get("exchange_ms_per_step"))
except Exception as e: print("FAILED", e)
PY
done
