#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the *reference* CUDA rasterizer and simple-knn (unmodified algorithm) for sm_100
# into oracle/_ref/, so the GPU parity tests and `bench.py --impl reference` can run them on the GPU box.
#
# Recipe:
#   * sources are read where they lie under /root/reference/submodules/{diff-gaussian-rasterization,simple-knn};
#     they are copied to a scratch dir under /tmp (the reference tree is read-only), never into this repo;
#   * two one-line patches are applied to the scratch copy:
#       - cuda_rasterizer/auxiliary.h:154   `p_view.z <= 0.2f` -> `0.001f`   (DAS3R README.md:41-44 tells the user
#         to make exactly this edit before `pip install`);
#       - cuda_rasterizer/rasterizer_impl.h gets `#include <cstdint>` (gcc 13 no longer pulls it in transitively);
#   * compiled like the reference's setup.py does (no fast-math; nvcc default -fmad=true), arch sm_100,
#     module names ref_dgr_C / ref_knn_C so they coexist with the new extension;
#   * outputs ONLY into oracle/_ref/ (git-ignored, not gpurun-ignored: the .so files travel to the GPU box).
# If /root/reference is absent (the GPU box), this script does nothing: the prebuilt files are used.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF=${DAS3R_REFERENCE:-/root/reference}/submodules
OUT="$HERE/_ref"
if [ ! -d "$REF/diff-gaussian-rasterization" ]; then
  echo "[build_ref] $REF not present - using prebuilt files in $OUT (if any)"; exit 0
fi
mkdir -p "$OUT"
STAMP="$OUT/.built"
if [ -f "$STAMP" ] && [ -f "$OUT/ref_dgr_C.so" ] && [ -f "$OUT/ref_knn_C.so" ] && [ "${1:-}" != "--force" ]; then
  echo "[build_ref] up to date"; exit 0
fi
TMP=$(mktemp -d /tmp/das3r_ref_build.XXXXXX)
trap 'rm -rf "$TMP"' EXIT
cp -r "$REF/diff-gaussian-rasterization" "$TMP/dgr"
cp -r "$REF/simple-knn" "$TMP/knn"
chmod -R u+w "$TMP"
sed -i 's/p_view\.z <= 0\.2f/p_view.z <= 0.001f/' "$TMP/dgr/cuda_rasterizer/auxiliary.h"
grep -q '0.001f' "$TMP/dgr/cuda_rasterizer/auxiliary.h" || { echo "near-plane patch did not apply"; exit 1; }
sed -i 's/#include <iostream>/#include <cstdint>\n#include <iostream>/' "$TMP/dgr/cuda_rasterizer/rasterizer_impl.h"

PY=${PYTHON:-python}
read -r TORCH_INC TORCH_LIB PY_INC EXT_SUFFIX <<<"$($PY - <<'PYEOF'
import sysconfig, os, torch
from torch.utils import cpp_extension as c
inc = " ".join("-I" + p for p in c.include_paths())
print(inc.replace(" ", ","), os.path.join(os.path.dirname(torch.__file__), "lib"), sysconfig.get_paths()["include"], sysconfig.get_config_var("EXT_SUFFIX"))
PYEOF
)"
TORCH_INC=${TORCH_INC//,/ }
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100,code=sm_100"
COMMON="-O3 -std=c++17 -D_GLIBCXX_USE_CXX11_ABI=1 -DTORCH_API_INCLUDE_EXTENSION_H $TORCH_INC -I$PY_INC"
NVFLAGS="$ARCH $COMMON --expt-relaxed-constexpr -Xcompiler -fPIC -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__"

build_dgr() {
  cd "$TMP/dgr"
  local D="-DTORCH_EXTENSION_NAME=ref_dgr_C -Ithird_party/glm"
  $NVCC $NVFLAGS $D -c cuda_rasterizer/rasterizer_impl.cu -o rasterizer_impl.o &
  $NVCC $NVFLAGS $D -c cuda_rasterizer/forward.cu -o forward.o &
  $NVCC $NVFLAGS $D -c cuda_rasterizer/backward.cu -o backward.o &
  $NVCC $NVFLAGS $D -c rasterize_points.cu -o rasterize_points.o &
  g++ $COMMON $D -fPIC -I/usr/local/cuda/include -c ext.cpp -o ext.o &
  wait
  g++ -shared -o "$OUT/ref_dgr_C.so" rasterizer_impl.o forward.o backward.o rasterize_points.o ext.o \
      -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_cuda -ltorch_python -lc10 -lc10_cuda \
      -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_LIB"
}
build_knn() {
  cd "$TMP/knn"
  local D="-DTORCH_EXTENSION_NAME=ref_knn_C"
  $NVCC $NVFLAGS $D -c simple_knn.cu -o simple_knn.o &
  $NVCC $NVFLAGS $D -c spatial.cu -o spatial.o &
  g++ $COMMON $D -fPIC -I/usr/local/cuda/include -c ext.cpp -o ext.o &
  wait
  g++ -shared -o "$OUT/ref_knn_C.so" simple_knn.o spatial.o ext.o \
      -L"$TORCH_LIB" -ltorch -ltorch_cpu -ltorch_cuda -ltorch_python -lc10 -lc10_cuda \
      -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_LIB"
}
( build_dgr ) &
( build_knn ) &
wait
date -u > "$STAMP"
ls -la "$OUT"
echo "[build_ref] done"
