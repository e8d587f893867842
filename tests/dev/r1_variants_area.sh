#!/usr/bin/env bash
# usage: r1_variants_area.sh <tag> <lib1> <lib2> ...  — config 3 bench line per library variant (development only)
TAG=$1; shift
mkdir -p gpurun_out
P=area-diameter-algorithm_b200
cp $P/libadk.so /tmp/libadk_orig.so
for v in "$@"; do
  cp $P/$v $P/libadk.so
  ADK_DEBUG=1 timeout 200 python bench.py --workload area_n1000 --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_${v}.json 2> gpurun_out/${TAG}_${v}.err
  echo "$v: $(cut -c1-75 gpurun_out/${TAG}_${v}.json)"; grep "adk\] rows" gpurun_out/${TAG}_${v}.err | tail -2
done
cp /tmp/libadk_orig.so $P/libadk.so
