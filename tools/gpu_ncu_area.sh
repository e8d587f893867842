#!/usr/bin/env bash
# ncu --set full captures of k_area_rows and k_area_prep on BASELINE config 3 (n = 1000), workspace capped so that ncu's
# save/restore of the table slots between its replay passes stays small.  usage: tools/gpu_ncu_area.sh <tag> [budget GB]
TAG=${1:-pa}; GB=${2:-40}
mkdir -p gpurun_out
CMD="python bench.py --workload area_n1000 --steps 1 --warmup 1 --no-cpu --memory-budget-gb $GB"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
cut -c1-200 gpurun_out/${TAG}_plain.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_area_rows -c 1 -f -o gpurun_out/${TAG}_rows $CMD > gpurun_out/${TAG}_ncu_rows.log 2>&1
echo "rows exit $?"; tail -2 gpurun_out/${TAG}_ncu_rows.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_area_prep -c 1 -f -o gpurun_out/${TAG}_prep $CMD > gpurun_out/${TAG}_ncu_prep.log 2>&1
echo "prep exit $?"; tail -2 gpurun_out/${TAG}_ncu_prep.log
ls -la gpurun_out/${TAG}_*.ncu-rep
