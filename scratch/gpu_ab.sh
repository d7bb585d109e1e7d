#!/bin/bash
# A/B bench of the variants in scratch/variants: scratch/gpu_ab.sh tag name1 name2 ...
TAG=$1; shift
mkdir -p gpurun_out
for n in "$@"; do
  FISHER_B200_LIB=$PWD/scratch/variants/lib_$n.so timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_$n.json"))
    print("$n value",round(d["value"],1),"argmax",d["argmax"], " ".join(f"{k[2:8]}={v['ms']/v['launches']*1e3:.0f}" for k,v in d["roofline"]["kernels"].items()))
except Exception as e:
    print("$n failed",e); print(open("gpurun_out/${TAG}_$n.err").read()[-1500:])
PY
done
