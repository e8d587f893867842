#!/usr/bin/env bash
# usage: r1_variants.sh <tag> <lib1> <lib2> ...  — short bench line per library variant (development only)
TAG=$1; shift
mkdir -p gpurun_out
P=area-diameter-algorithm_b200
cp $P/libadk.so /tmp/libadk_orig.so
for v in "$@"; do
  cp $P/$v $P/libadk.so
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${TAG}_${v}.json 2> gpurun_out/${TAG}_${v}.err
  echo "$v: $(cut -c1-75 gpurun_out/${TAG}_${v}.json)"
done
cp /tmp/libadk_orig.so $P/libadk.so
