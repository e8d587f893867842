#!/usr/bin/env bash
# ncu --set full capture of one k_ad_cells launch (the D = 6 solve of the first sweep).  usage: tools/gpu_ncu_cells.sh <tag>
TAG=${1:-p}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_ad_cells -s 4 -c 1 -f -o gpurun_out/${TAG}_cells $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
