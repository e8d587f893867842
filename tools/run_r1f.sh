mkdir -p gpurun_out
timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 3 > gpurun_out/r1f_bench_area_n1000_n1.json 2> gpurun_out/r1f_area.err; echo "area rc=$?"
timeout 360 python -m pytest tests -m gpu -x -q > gpurun_out/r1f_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r1f_pytest_gpu.log
timeout 200 python bench.py > gpurun_out/r1f_bench_n1.json 2> gpurun_out/r1f_bench.err; echo "bench rc=$?"
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1f_launches_area_n1000.csv python bench.py --workload area_n1000 --steps 1 --warmup 3 --no-cpu > gpurun_out/r1f_ncu.log 2>&1; echo "ncu rc=$?"
