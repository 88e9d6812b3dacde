#!/bin/bash
# Sensitivity of the coarse touch legs (config 2 on G-wt2, bench.py's G-org leg) to the swept-window knobs.
run() {
  echo "== $*"
  env "$@" timeout 600 python scripts/bench_configs.py --legs touch --scale 1.0 2>/dev/null | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['leg'], '%.4g' % d['segments_per_s'])"
  env "$@" timeout 600 python scripts/touch_probe.py 2>&1 | tail -1
}
run OSR_SWEEP_SHORT_TILE=8
run OSR_SWEEP_SHORT_TILE=4
run OSR_SWEEP_SHORT_TILE=6
run OSR_SWEEP_SHORT_TILE=12
run OSR_SWEEP_SHORT_TILE=16
run OSR_SWEEP_RSTART=0.0025
run OSR_SWEEP_RSTART=0.01
run OSR_SWEEP_RMAX=0.01
run OSR_SWEEP_RMAX=0.04
