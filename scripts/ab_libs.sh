#!/bin/bash
# A/B of development builds (build.py --variant TAG ...): parity tests + segment legs per library, same box, same run.
# usage: scripts/ab_libs.sh TAG...   ("base" = the product library)
for v in "$@"; do
  if [ "$v" != base ]; then export OSR_LIB=$PWD/puzzletunneldiscovery_b200/libosrgpu_$v.so; else unset OSR_LIB; fi
  echo "== variant '$v'"
  timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
  timeout 900 python scripts/bench_configs.py --legs ${LEGS:-touch,touch_fine,clearance,edges,ompl_motions,contacts} --scale 1.0 2>/dev/null | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['leg'], '%.4g' % d.get('segments_per_s', d.get('edges_per_s', d.get('motions_per_s', d.get('contacts_per_s', 0)))))
"
  timeout 600 python scripts/touch_probe.py 2>&1 | tail -1
done
