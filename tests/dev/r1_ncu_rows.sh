#!/usr/bin/env bash
# usage: r1_ncu_rows.sh <tag>   — plain run, then one full capture of the row-sweep DP kernel (config 3, n = 1000, heavy batch)
TAG=${1:-a2}
mkdir -p gpurun_out
CMD="python bench.py --workload area_n1000 --steps 1 --warmup 1 --no-cpu"
timeout 200 $CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_area_rows -s 2 -c 1 -f -o gpurun_out/${TAG}_prof_area_rows $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -n 2 gpurun_out/${TAG}_plain.log gpurun_out/${TAG}_ncu2.log | cut -c1-300
