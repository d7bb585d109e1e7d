#!/bin/bash
# quick visit: gpu tests + bench (no ncu).  usage: scratch/gpu_quick.sh <tag> [pytest-args]
TAG=${1:-q}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench.json"))
    print("value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"argmax",d["argmax"])
    for k,v in d["roofline"]["kernels"].items(): print(f"  {k:18s} {v['ms']/v['launches']*1e3:8.1f} us/launch share {v['share']:.3f}")
except Exception as e:
    print("bench parse failed",e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
