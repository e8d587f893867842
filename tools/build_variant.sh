#!/usr/bin/env bash
# development: build libadk variants with extra -D flags into build/variants/<name>.so   usage: build_variant.sh <name> [-DX=Y ...]
set -e
NAME=$1; shift
mkdir -p build/variants
cd area-diameter-algorithm_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -shared "$@" adk_capi.cu adk_precompute.cu adk_area.cu adk_ad.cu adk_multi.cu adk_tiled.cu -ldl -lpthread -o ../../build/variants/$NAME.so 2>&1 | grep -i "error" || true
