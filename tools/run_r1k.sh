mkdir -p gpurun_out
ADK_DEBUG=1 timeout 120 python bench.py --workload area_n1000 --steps 1 --warmup 1 --no-cpu > gpurun_out/r1k_area_debug.json 2> gpurun_out/r1k_area_debug.err; echo "rc=$?"
grep "rows kernel" gpurun_out/r1k_area_debug.err | tail -n 3
