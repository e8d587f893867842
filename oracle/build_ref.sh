#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the UNMODIFIED reference solver into oracle/_ref/.
#
# Compiles /root/reference/src/common/*.cpp where they lie (nothing is copied
# into this repo) plus oracle/ref_shim.cpp (our extern "C" face) with plain g++.
# The reference's own CMake build is NOT used: it cannot configure here because
# external/json is an empty submodule, and Parser.cpp (the only user of that
# dependency) is not on the solver path, so it is simply left out.
#
# Outputs (git-ignored, but shipped to the GPU box by gpurun):
#   oracle/_ref/libadkref.so      native build, long double = x87 80-bit  (THE oracle)
#   oracle/_ref/libadkref_f64.so  same sources, -mlong-double-64          (binary64 variant)
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
ref="${ADK_REFERENCE_ROOT:-/root/reference}"
src="$ref/src/common"
out="$here/_ref"
if [ ! -d "$src" ]; then
    echo "build_ref.sh: $src not present; keeping prebuilt files in $out" >&2
    exit 0
fi
mkdir -p "$out"
files=("$src/CustomMath.cpp" "$src/Utils.cpp" "$src/Eppstein.cpp" "$src/Antipodal.cpp" "$src/AntipodalOptimized.cpp")
common=(-std=c++17 -O3 -fPIC -shared -w -include "$here/ref_compat.h" -I"$src")
g++ "${common[@]}" "$here/ref_shim.cpp" "${files[@]}" -o "$out/libadkref.so" &
g++ "${common[@]}" -mlong-double-64 -DADK_REF_F64 "$here/ref_shim.cpp" "${files[@]}" -o "$out/libadkref_f64.so" &
wait
echo "built $out/libadkref.so $out/libadkref_f64.so"
