#!/usr/bin/env bash
# Development session on the GPU box: parity suite, bench (own arm + reference arm), ncu launch list and one full
# capture of the top kernel.  usage: r1_gpu_session.sh <tag> [skip]
TAG=${1:-r1b}; SKIP=${2:-4}
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log)
(timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?" >> gpurun_out/${TAG}_bench.err)
(timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err)
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ad_cells -s $SKIP -c 1 -f -o gpurun_out/${TAG}_prof_cells $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
$CMD > gpurun_out/${TAG}_plain3.log 2>&1 &&
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_ad_cells -s 5 -c 5 --csv --log-file gpurun_out/${TAG}_traffic.csv $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
tail -3 gpurun_out/${TAG}_pytest.log; cut -c1-900 gpurun_out/${TAG}_bench.json; tail -n 2 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log
