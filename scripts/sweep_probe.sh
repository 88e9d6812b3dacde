#!/bin/bash
# A/B of the swept-window parameters of K2 (OSR_SWEEP_TILES x OSR_SWEEP_RMAX) on the segment legs.
LEGS=${LEGS:-touch,touch_fine,clearance}
CFGS=${CFGS:-"0:0.02 64:0.005 64:0.01 64:0.02"}
for cfg in $CFGS; do
  t=${cfg%%:*}; r=${cfg##*:}
  echo "== tiles=$t rmax=$r"
  OSR_SWEEP_TILES=$t OSR_SWEEP_RMAX=$r python scripts/bench_configs.py --legs $LEGS 2>&1 | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['leg'], round(d.get('segments_per_s', d.get('edges_per_s', 0))), d.get('same_first_contact_step_as_oracle_on_prefix', d.get('labels_equal_oracle_on_prefix', d.get('max_abs_distance_error_vs_oracle_on_prefix'))))
"
done
