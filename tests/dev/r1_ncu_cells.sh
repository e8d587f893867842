#!/usr/bin/env bash
# usage: r1_ncu_cells.sh <tag> <skip>   — one full ncu capture of k_ad_cells (launch index <skip> of the bench command)
TAG=${1:-cX}; SKIP=${2:-4}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ad_cells -s $SKIP -c 1 -f -o gpurun_out/${TAG}_prof_cells $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
tail -n 2 gpurun_out/${TAG}_ncu2.log
