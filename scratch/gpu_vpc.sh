#!/bin/bash
for v in "$@"; do
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --views-per-chunk $v > gpurun_out/vpc_$v.json 2> gpurun_out/vpc_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/vpc_$v.json"))
    print("V=$v value",round(d["value"],1)," ".join(f"{k[2:8]}={v['ms']/v['launches']*1e3:.0f}" for k,v in d["roofline"]["kernels"].items()))
except Exception as e:
    print("V=$v failed",e); print(open("gpurun_out/vpc_$v.err").read()[-1500:])
PY
done
