#!/usr/bin/env bash
# Development session on the GPU box: parity suite, bench, ncu launch list and one full capture of the top kernel.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log)
(timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err)
(timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err)
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ad_dp -s 4 -c 1 -f -o gpurun_out/prof_ad_dp $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/pytest.log; cat gpurun_out/bench.json | cut -c1-600; tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log
