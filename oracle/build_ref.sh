#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the UNMODIFIED reference rasterizer as the bit-level CUDA oracle.
#
# Compiles thirdparty/diff-gaussian-rasterization-w-pose from the sources where they lie under
# /root/reference (nothing is copied into this repo; only build outputs land in oracle/_ref/).
# Two edits are applied to the scratch copy, both documented in SURVEY.md:
#   1. near plane 0.2f -> 0.001f in cuda_rasterizer/auxiliary.h:154 — exactly what Splat-SLAM's
#      own Dockerfile:20 / README.md:88-92 do at install time ("as-installed" reference);
#   2. `#include <cstdint>` in cuda_rasterizer/rasterizer_impl.h (GCC 13 needs it for uintptr_t).
# The module is renamed ref_dgr so it can be imported next to the new `diff_gaussian_rasterization`.
#
# Output: oracle/_ref/ref_dgr/{__init__.py,_C*.so}   (git-ignored, travels to the GPU box via gpurun)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}/thirdparty/diff-gaussian-rasterization-w-pose"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref: $REF not present (GPU box?) — using prebuilt oracle/_ref if any" >&2
  exit 0
fi
if ls "$OUT"/ref_dgr/_C*.so >/dev/null 2>&1 && [ "${FORCE:-0}" != "1" ]; then
  echo "build_ref: oracle/_ref/ref_dgr already built"
  exit 0
fi
TMP="$(mktemp -d /tmp/ref_dgr_build.XXXXXX)"
cp -r "$REF"/. "$TMP"/
chmod -R u+w "$TMP"
cd "$TMP"
sed -i 's/p_view\.z <= 0\.2f/p_view.z <= 0.001f/' cuda_rasterizer/auxiliary.h
grep -q 'p_view.z <= 0.001f' cuda_rasterizer/auxiliary.h
sed -i 's/#include <iostream>/#include <cstdint>\n#include <iostream>/' cuda_rasterizer/rasterizer_impl.h
mv diff_gaussian_rasterization ref_dgr
sed -i 's/diff_gaussian_rasterization/ref_dgr/g' setup.py
export TORCH_CUDA_ARCH_LIST="10.0"
export MAX_JOBS="${MAX_JOBS:-8}"
python setup.py build_ext --inplace > build.log 2>&1 || { tail -50 build.log; exit 1; }
mkdir -p "$OUT/ref_dgr"
cp ref_dgr/__init__.py "$OUT/ref_dgr/"
cp ref_dgr/_C*.so "$OUT/ref_dgr/"
echo "build_ref: built $(ls "$OUT"/ref_dgr)"
rm -rf "$TMP"
