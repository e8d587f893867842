#!/usr/bin/env bash
# Builds the reference's UNMODIFIED gtest (src/test/main.cpp), pybind module (src/py-binding/main.cpp) and google-benchmark
# driver (src/bench/main.cpp) against this
# repo's GPU-backed libadk_host.so instead of the reference's src/common solver sources.  Proves the drop-in boundary:
# same headers (the reference's own declarations), our definitions at link time.  Only Parser.cpp (JSON I/O, not on the
# solver path) is taken from the reference.  Needs /root/reference; outputs go to tests/dropin/_build (git-ignored).
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
root="$(cd "$here/../.." && pwd)"
ref="${ADK_REFERENCE_ROOT:-/root/reference}"
# a caller of the reference's public helpers (PointsInTriangle, ForAllAntipodalPairs, ComputeDiameter), written against the
# reference's header names and built against this repo's mirror; needs no reference sources
mkdir -p "$here/_build"
g++ -std=c++17 -O2 -I"$root/area-diameter-algorithm_b200/host" "$here/helpers_caller.cpp" -o "$here/_build/helpers_caller" \
    -L"$root/area-diameter-algorithm_b200" -ladk_host -ladk -Wl,-rpath,"$root/area-diameter-algorithm_b200" -Wl,-rpath,'$ORIGIN/../../../area-diameter-algorithm_b200'
[ -d "$ref/src/test" ] || { echo "build_dropin.sh: $ref not present, keeping prebuilt files" >&2; exit 0; }
out="$here/_build"; mkdir -p "$out"
json_inc="$(python - <<'PY'
import os, sys
for p in sys.path:
    q = os.path.join(p, "include", "cudnn_frontend", "thirdparty")
    if os.path.exists(os.path.join(q, "nlohmann", "json.hpp")):
        print(q); break
PY
)"
[ -n "$json_inc" ] || { echo "nlohmann/json.hpp not found" >&2; exit 1; }
gt="$ref/external/googletest/googletest"
pkg="$root/area-diameter-algorithm_b200"
common=(-std=c++17 -O2 -w -include "$here/dropin_compat.h" -I"$json_inc")
if [ ! -f "$out/gtest-all.o" ]; then
  g++ -std=c++17 -O1 -w -I"$gt/include" -I"$gt" -c "$gt/src/gtest-all.cc" -o "$out/gtest-all.o" &
  g++ -std=c++17 -O1 -w -I"$gt/include" -I"$gt" -c "$gt/src/gtest_main.cc" -o "$out/gtest_main.o" &
fi
g++ "${common[@]}" -c "$ref/src/common/Parser.cpp" -o "$out/Parser.o" &
g++ "${common[@]}" -I"$gt/include" -c "$ref/src/test/main.cpp" -o "$out/ref_test_main.o" &
wait
g++ "$out/ref_test_main.o" "$out/Parser.o" "$out/gtest-all.o" "$out/gtest_main.o" -o "$out/ref_test_dropin" \
    -L"$pkg" -ladk_host -ladk -Wl,-rpath,"$pkg" -Wl,-rpath,'$ORIGIN/../../../area-diameter-algorithm_b200' -lpthread
# the reference's pybind11 module `thesis`, same treatment
pyinc="$(python -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
pbinc="$(python -c 'import pybind11; print(pybind11.get_include())')"
ext="$(python -c 'import sysconfig; print(sysconfig.get_config_var("EXT_SUFFIX"))')"
g++ "${common[@]}" -fPIC -shared -I"$pyinc" -I"$pbinc" "$ref/src/py-binding/main.cpp" -o "$out/thesis$ext" \
    -L"$pkg" -ladk_host -ladk -Wl,-rpath,"$pkg" -Wl,-rpath,'$ORIGIN/../../../area-diameter-algorithm_b200'
# the reference's google-benchmark driver (src/bench/main.cpp), same treatment; google/benchmark itself is compiled from
# the sources vendored by the reference (external/benchmark/src/*.cc, plain g++, no cmake)
gb="$ref/external/benchmark"
if [ -d "$gb/src" ]; then
  mkdir -p "$out/gbench"
  if [ ! -f "$out/gbench/benchmark.o" ]; then
    for f in "$gb"/src/*.cc; do
      b="$(basename "$f" .cc)"; [ "$b" = benchmark_main ] && continue
      g++ -std=c++17 -O1 -w -I"$gb/include" -I"$gb/src" -DHAVE_STD_REGEX -DHAVE_STEADY_CLOCK -DNDEBUG -DBENCHMARK_STATIC_DEFINE -c "$f" -o "$out/gbench/$b.o" &
    done
    wait
  fi
  g++ "${common[@]}" -I"$gb/include" -DBENCHMARK_STATIC_DEFINE -c "$ref/src/bench/main.cpp" -o "$out/ref_bench_main.o"
  g++ "$out/ref_bench_main.o" "$out/Parser.o" "$out"/gbench/*.o -o "$out/ref_bench_dropin" \
      -L"$pkg" -ladk_host -ladk -Wl,-rpath,"$pkg" -Wl,-rpath,'$ORIGIN/../../../area-diameter-algorithm_b200' -lpthread
fi
echo "built $out/ref_test_dropin $out/thesis$ext $out/ref_bench_dropin"
