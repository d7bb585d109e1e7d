#!/bin/bash
for p in 0 1 0 1; do
  FB_LANE_PRIORITY=$p python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('prio $p value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'argmax', d['argmax'])"
done
