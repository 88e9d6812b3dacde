#!/bin/bash
# A/B of one swept-window knob of K2 on the segment legs:  VAR=OSR_SWEEP_SHORT_TILE VALS="32 16 8 4" scripts/sweep_probe.sh
LEGS=${LEGS:-touch,touch_fine,clearance}
VAR=${VAR:-OSR_SWEEP_SHORT_TILE}
for v in ${VALS:-32 16 8 4}; do
  echo "== $VAR=$v"
  env $VAR=$v python scripts/bench_configs.py --legs $LEGS 2>&1 | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['leg'], round(d.get('segments_per_s', d.get('edges_per_s', 0))), d.get('same_first_contact_step_as_oracle_on_prefix', d.get('labels_equal_oracle_on_prefix', d.get('max_abs_distance_error_vs_oracle_on_prefix'))))
"
  env $VAR=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bench touch (G-org)', round(d['touch']['segments_per_s']))"
done
