#!/usr/bin/env bash
# Build the UNMODIFIED reference CUDA rasterizer and simple-knn, in place from
# /root/reference, into oracle/_ref/*.so (git-ignored, travels to the GPU box).
# Test infrastructure only: used as the golden in `-m gpu` parity tests, to
# generate tests/golden/ fixtures, and as bench.py's reference arm.
# The reference's own build system (setup.py / CMake) is not used.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${SPLATLOC_REFERENCE:-/root/reference}"
DGR="$REF/submodules/diff-gaussian-rasterization"
KNN="$REF/submodules/simple-knn"
OUT="$HERE/_ref"
if [ ! -d "$DGR/cuda_rasterizer" ]; then
  echo "build_ref: $DGR not present; keeping any prebuilt oracle/_ref" >&2
  exit 0
fi
mkdir -p "$OUT"
NVCC="${NVCC:-nvcc}"
# same flags as DGR/setup.py:31-36 (-O3, glm include) + the sm_100a target;
# `--pre-include cstdint`: rasterizer_impl.h:25,39-61 relies on a transitive
# <cstdint> that GCC 13 no longer provides.
COMMON=(-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC
        -Xcompiler -fno-gnu-unique --pre-include cstdint -w)
build_raster() {  # $1 = output name, $2.. = extra flags
  local name="$1"; shift
  if [ "$OUT/$name" -nt "$DGR/cuda_rasterizer/backward.cu" ] && [ "$OUT/$name" -nt "$HERE/ref_shim.cu" ]; then
    echo "build_ref: $name up to date"; return
  fi
  "$NVCC" "${COMMON[@]}" "$@" -DREF_SHIM_RASTER -I"$DGR" -I"$DGR/third_party/glm" -shared \
      "$HERE/ref_shim.cu" "$DGR/cuda_rasterizer/forward.cu" "$DGR/cuda_rasterizer/backward.cu" \
      "$DGR/cuda_rasterizer/rasterizer_impl.cu" -o "$OUT/$name"
  echo "build_ref: built $name"
}
build_raster libref_raster_c4.so &
build_raster libref_raster_c3.so --pre-include "$HERE/ref_config_c3.h" &
if [ ! "$OUT/libref_knn.so" -nt "$KNN/simple_knn.cu" ] || [ ! "$OUT/libref_knn.so" -nt "$HERE/ref_shim.cu" ]; then
  "$NVCC" "${COMMON[@]}" -DREF_SHIM_KNN -I"$KNN" -shared "$HERE/ref_shim.cu" "$KNN/simple_knn.cu" \
      -o "$OUT/libref_knn.so" && echo "build_ref: built libref_knn.so" &
fi
wait
ls -la "$OUT"
