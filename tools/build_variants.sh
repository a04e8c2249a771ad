#!/bin/bash
# build alternative libdca_b200.so files into tools/variants/ (benched on the GPU box by tools/gpu_variants.sh)
# usage: tools/build_variants.sh name1="-DFLAG=.." name2="..."
cd "$(dirname "$0")/.."
mkdir -p tools/variants
rm -f tools/variants/*.so
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -shared -Xcompiler -fPIC $flags \
       -o tools/variants/$name.so hmp2g_b200/csrc/dca_b200.cu &
done
wait
ls -la tools/variants
