#!/usr/bin/env bash
# configs 3 and 5 on one 8-GPU box.  usage: tools/gpu_scale_aux.sh <tag>
TAG=${1:-r2}
mkdir -p gpurun_out
run() { # workload N extra
  W=$1; N=$2; shift 2
  if [ "$N" = 1 ]; then CMD="python bench.py"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 296$N bench.py"; fi
  timeout 600 $CMD --workload $W --gpus $N --steps 3 --warmup 2 --no-cpu "$@" > gpurun_out/${TAG}_${W}_n$N.json 2> gpurun_out/${TAG}_${W}_n$N.err
  python - $W $N gpurun_out/${TAG}_${W}_n$N.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[3]).read().strip().splitlines()[-1])
    print(sys.argv[1], "N", sys.argv[2], "ms/step %.1f" % d["ms_per_step"], "value %.4g" % d["value"], "kernel %.1f ms" % d["roofline"]["avg_launch_ms"], "share %.2f" % d["roofline"]["share_of_step"])
except Exception as e: print(sys.argv[1], sys.argv[2], "FAILED", e)
PY
}
for N in 8 4 2; do run area_n1000 $N; done
run cells_batch 4
