#!/bin/bash
# GPU-box visit used during development: K2 tile kernel occupancy variants.
mkdir -p gpurun_out
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["touch"]["segments_per_s"], d["touch"]["validity_checks_per_s"])'
for b in 2 3 4; do
  echo "transit blocks $b"; OSR_TRANSIT_BLOCKS=$b timeout 300 python bench.py --no-cpu-baseline --steps 5 --warmup 3 | python -c "$P"
  OSR_TRANSIT_BLOCKS=$b timeout 600 python scripts/bench_configs.py --scale 0.25 --legs touch,touch_fine,edges 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.strip()); continue
    print(d['leg'], {k: v for k, v in d.items() if k.endswith('_per_s') or 'oracle' in k})"
done
OSR_TRANSIT_BLOCKS=3 timeout 600 python -m pytest tests -m gpu -x -q -k "segment or edge or visib" 2>&1 | tail -2
