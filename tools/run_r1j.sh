mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --workload area_n1000 --steps 3 --warmup 3 > gpurun_out/r1j_bench_area_n1000_n8.json 2> gpurun_out/r1j_area_n8.err; echo "area n8 rc=$?"
tail -n 1 gpurun_out/r1j_bench_area_n1000_n8.json | cut -c1-400
