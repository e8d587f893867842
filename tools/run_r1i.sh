mkdir -p gpurun_out
timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 3 --no-cpu > gpurun_out/r1i_bench_area_n1000_n1.json 2> gpurun_out/r1i_area.err; echo "area rc=$?"
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r1i_pytest_gpu.log
