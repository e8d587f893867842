#!/usr/bin/env bash
# round-end evidence on one B200: headline bench + reference arm, ncu launch list and DRAM traffic of the same command,
# config 3 bench + ncu --set full of k_area_rows.  usage: tools/gpu_final.sh <tag>
TAG=${1:-r2f}
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref exit $?"
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list exit $?"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_ad_cells -s 5 -c 5 --csv --log-file gpurun_out/${TAG}_traffic_cells.csv $CMD > gpurun_out/${TAG}_ncu_traffic.log 2>&1
echo "traffic exit $?"
python bench.py --workload area_n1000 > gpurun_out/${TAG}_bench_area_n1000_n1.json 2> gpurun_out/${TAG}_bench_area_n1000_n1.err; echo "area exit $?"
ACMD="python bench.py --workload area_n1000 --steps 1 --warmup 1 --no-cpu --memory-budget-gb 90"
$ACMD > gpurun_out/${TAG}_area_plain.json 2> gpurun_out/${TAG}_area_plain.err &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_area_rows -c 1 -f -o gpurun_out/${TAG}_rows $ACMD > gpurun_out/${TAG}_ncu_rows.log 2>&1
echo "rows ncu exit $?"; tail -2 gpurun_out/${TAG}_ncu_rows.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_area.csv $ACMD > gpurun_out/${TAG}_ncu_launches_area.log 2>&1
echo "area launch list exit $?"
cut -c1-300 gpurun_out/${TAG}_bench_n1.json; cut -c1-300 gpurun_out/${TAG}_bench_reference.json; cut -c1-300 gpurun_out/${TAG}_bench_area_n1000_n1.json; cut -c1-200 gpurun_out/${TAG}_area_plain.json
