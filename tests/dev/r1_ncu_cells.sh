#!/usr/bin/env bash
# usage: r1_ncu_cells.sh <tag> <skip>   — bench line, launch list and one full capture of k_ad_cells (sigma=2, D=4 launch)
TAG=${1:-cX}; SKIP=${2:-2}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ad_cells -s $SKIP -c 1 -f -o gpurun_out/${TAG}_prof_cells $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -n 2 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log
