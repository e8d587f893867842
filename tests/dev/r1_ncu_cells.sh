#!/usr/bin/env bash
mkdir -p gpurun_out
(timeout 300 python bench.py --no-cpu > gpurun_out/c6_bench.json 2> gpurun_out/c6_bench.err; echo "bench exit $?" >> gpurun_out/c6_bench.err)
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c6_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ad_cells -s 4 -c 1 -f -o gpurun_out/c6_prof_cells $CMD > gpurun_out/ncu2.log 2>&1
cut -c1-300 gpurun_out/c6_bench.json; tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log
