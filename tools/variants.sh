#!/bin/bash
# Developer tool: build variants of libdas3r_raster.so with different -D tunables into var/ (git-ignored, travels with
# gpurun). Usage: tools/variants.sh name "-DDAS3R_FWD_K=8 -DDAS3R_BWD_K=8" [name2 "flags2" ...]
# On the GPU box: cp var/<name>.so das3r_b200/libdas3r_raster.so before running a tool.
set -e
cd "$(dirname "$0")/.."
python das3r_b200/build.py > /dev/null
mkdir -p var /tmp/var_build
cp das3r_b200/libdas3r_raster.so var/base.so
ARCH="-gencode arch=compute_100a,code=sm_100a"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  for f in forward backward; do
    /usr/local/cuda/bin/nvcc $ARCH -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -c das3r_b200/csrc/$f.cu -o /tmp/var_build/${name}_$f.o &
  done
  wait
  /usr/local/cuda/bin/nvcc $ARCH -shared -o var/$name.so /tmp/var_build/${name}_forward.o /tmp/var_build/${name}_backward.o \
      $(ls das3r_b200/build/*.o | grep -v -e /forward.o -e /backward.o) -cudart static
  echo built var/$name.so
done
