#!/bin/bash
# A/B run of kernel variants on the GPU box: one short bench per library in build_variants/ (or the ones named).
# usage: tools/gpu_ab.sh <tag> [variant names...]; writes gpurun_out/<tag>_ab.jsonl (one bench line per variant)
TAG=${1:-ab}; shift
mkdir -p gpurun_out
names="$@"; [ -z "$names" ] && names=$(ls build_variants | sed 's/^lib_//; s/\.so$//')
for n in $names; do
  FISHER_B200_LIB=$PWD/build_variants/lib_$n.so timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err
  python - "$n" gpurun_out/${TAG}_$n.json <<'PY' | tee -a gpurun_out/${TAG}_ab.txt
import json, sys
n, f = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    k = d["roofline"]["kernels"]
    per = {kk: round(v["ms"] / v["launches"] * 1e3) for kk, v in k.items()}
    print(f"{n:14s} value {d['value']:.0f} e2e {d['e2e']['value']:.0f} argmax {d['argmax']} us/launch {per}")
except Exception as e:
    print(n, "FAILED", e)
PY
done
