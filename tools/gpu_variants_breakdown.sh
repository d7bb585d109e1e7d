#!/bin/bash
# per-kernel times (tools/kernel_breakdown.py) of the built library and of every build_variants/lib_*.so at 3M and 500k Gaussians
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/${1:-variants}_breakdown.txt
: > $out
for lib in fisher_nerf_customized_b200/libfisher_b200.so build_variants/lib_*.so; do
  [ -f $lib ] || continue
  v=$(basename $lib .so)
  echo -n "$v 3M: " >> $out
  FISHER_B200_LIB=$lib timeout 300 python tools/kernel_breakdown.py 3000000 128 4 30 30 >> $out 2>&1
  echo -n "$v 500k: " >> $out
  FISHER_B200_LIB=$lib timeout 300 python tools/kernel_breakdown.py 500000 256 4 >> $out 2>&1
done
cat $out
