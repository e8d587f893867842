mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r1g_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r1g_pytest_gpu.log
timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 3 > gpurun_out/r1g_bench_area_n1000_n1.json 2> gpurun_out/r1g_area.err; echo "area rc=$?"
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1g_launches_area_n1000.csv python bench.py --workload area_n1000 --steps 1 --warmup 3 --no-cpu > gpurun_out/r1g_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 500 ncu --set full --clock-control none --import-source on -k regex:k_area_rows -c 1 -f -o gpurun_out/r1g_k_area_rows python bench.py --workload area_n1000 --steps 1 --warmup 0 --no-cpu > gpurun_out/r1g_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/r1g_ncu_full.log
