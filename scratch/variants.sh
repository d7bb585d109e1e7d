#!/bin/bash
# build kernel variants: scratch/variants.sh name "-DFLAG=1 ..." [name2 "flags2" ...]
mkdir -p scratch/variants
while [ $# -ge 2 ]; do
  n=$1; f=$2; shift 2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $f \
     -o scratch/variants/lib_$n.so fisher_nerf_customized_b200/csrc/fb_api.cu &
done
wait; ls -la scratch/variants
