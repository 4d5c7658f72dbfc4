#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the UNMODIFIED reference simple-knn extension (distCUDA2) as the bit-level oracle
# for splat_raster_knn_mean_dist2.
#
# Compiles thirdparty/simple-knn from the sources where they lie under /root/reference (nothing is copied into this
# repo; only build outputs land in oracle/_ref/). One edit is applied to the scratch copy: `#include <cfloat>` at
# the top of simple_knn.cu (FLT_MAX is otherwise undeclared with this toolchain). The module is renamed
# ref_simple_knn.
#
# Output: oracle/_ref/ref_simple_knn/_C*.so   (git-ignored, travels to the GPU box via gpurun)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}/thirdparty/simple-knn"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref_knn: $REF not present (GPU box?) — using prebuilt oracle/_ref if any" >&2
  exit 0
fi
if ls "$OUT"/ref_simple_knn/_C*.so >/dev/null 2>&1 && [ "${FORCE:-0}" != "1" ]; then
  echo "build_ref_knn: oracle/_ref/ref_simple_knn already built"
  exit 0
fi
TMP="$(mktemp -d /tmp/ref_knn_build.XXXXXX)"
cp -r "$REF"/. "$TMP"/
chmod -R u+w "$TMP"
cd "$TMP"
sed -i '0,/#include "cuda_runtime.h"/s//#include <cfloat>\n#include "cuda_runtime.h"/' simple_knn.cu
rm -rf simple_knn && mkdir ref_simple_knn && touch ref_simple_knn/__init__.py
sed -i 's/simple_knn\._C/ref_simple_knn._C/; s/name="simple_knn"/name="ref_simple_knn"/' setup.py
export TORCH_CUDA_ARCH_LIST="10.0"
export MAX_JOBS="${MAX_JOBS:-8}"
python setup.py build_ext --inplace > build.log 2>&1 || { tail -50 build.log; exit 1; }
mkdir -p "$OUT/ref_simple_knn"
cp ref_simple_knn/__init__.py "$OUT/ref_simple_knn/"
cp ref_simple_knn/_C*.so "$OUT/ref_simple_knn/"
echo "build_ref_knn: built $(ls "$OUT"/ref_simple_knn)"
rm -rf "$TMP"
