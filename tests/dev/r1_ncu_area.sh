#!/usr/bin/env bash
# usage: r1_ncu_area.sh <tag>   — launch list and one full capture of the area path's DP kernel (config 3, n = 1000)
TAG=${1:-a1}
mkdir -p gpurun_out
CMD="python bench.py --workload area_n1000 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_area_dp -s 1 -c 1 -f -o gpurun_out/${TAG}_prof_area_dp $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -n 2 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log
