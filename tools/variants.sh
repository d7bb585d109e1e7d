#!/bin/bash
# Build named kernel variants of libfisher_b200.so into build_variants/ for A/B runs on the GPU box.
# usage: tools/variants.sh name1="-DFLAG=.. -DFLAG2=.." name2="..."
set -e
cd "$(dirname "$0")/.."
mkdir -p build_variants
for spec in "$@"; do
  name="${spec%%=*}"; flags="${spec#*=}"
  [ "$name" = "$spec" ] && flags=""
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
       -o build_variants/lib_$name.so fisher_nerf_customized_b200/csrc/fb_api.cu 2>&1 | grep -i "error" || true
  echo "built build_variants/lib_$name.so ($flags)"
done
