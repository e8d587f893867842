mkdir -p gpurun_out
timeout 200 python bench.py > gpurun_out/r1h_bench_n1.json 2> gpurun_out/r1h_bench.err; echo "bench rc=$?"
timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 3 > gpurun_out/r1h_bench_area_n1000_n1.json 2> gpurun_out/r1h_area.err; echo "area rc=$?"
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1h_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/r1h_smoke.log
timeout 240 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1h_bench_reference.json 2> gpurun_out/r1h_ref.err; echo "ref rc=$?"
