#!/bin/bash
# A/B of the two chunk lanes: e2e value (profiling off => lanes active) with lanes on/off
python - <<'PY'
import json, subprocess, os, sys
for lanes in (1, 2, 1, 2):
    env = dict(os.environ, FB_CHUNK_LANES=str(lanes))
    out = subprocess.run([sys.executable, "bench.py", "--steps", "5", "--warmup", "3", "--no-cpu-baseline"], env=env, capture_output=True, text=True)
    try:
        d = json.loads(out.stdout.strip().splitlines()[-1])
        print("lanes", lanes, "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "argmax", d["argmax"])
    except Exception as e:
        print("failed", e, out.stderr[-2000:])
PY
