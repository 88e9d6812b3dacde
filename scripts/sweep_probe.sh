#!/bin/bash
# Sensitivity of the segment legs to the swept-window knobs (one setting per line: "VAR=value ...").
run() {
  echo "== $*"
  env "$@" timeout 600 python scripts/bench_configs.py --legs ${LEGS:-touch,touch_fine,edges,clearance} --scale 1.0 2>/dev/null | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['leg'], '%.4g' % d.get('segments_per_s', d.get('edges_per_s', 0)))"
  env "$@" timeout 600 python scripts/touch_probe.py 2>&1 | tail -1
}
while read -r line; do [ -n "$line" ] && run $line; done <<< "${SETTINGS:-OSR_SWEEP_NB0=4
OSR_SWEEP_NB0=8
OSR_SWEEP_NB0=16
OSR_SWEEP_NB0=2}"
