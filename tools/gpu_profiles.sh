#!/usr/bin/env bash
# round-2 evidence for profiles/: bench lines, the ncu launch list of the same command, DRAM traffic of the dominant kernel.
TAG=${1:-r2}
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref exit $?"
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list exit $?"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_ad_cells -s 5 -c 5 --csv --log-file gpurun_out/${TAG}_traffic_cells.csv $CMD > gpurun_out/${TAG}_ncu_traffic.log 2>&1
echo "traffic exit $?"
[ -x tools/issue_peaks ] && tools/issue_peaks > gpurun_out/${TAG}_issue_peaks.json 2>&1
cut -c1-300 gpurun_out/${TAG}_bench_n1.json; cut -c1-300 gpurun_out/${TAG}_bench_reference.json
