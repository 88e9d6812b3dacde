#!/bin/bash
# A/B of K2's kept-pair windows: OSR_SWEEP_RESOLVE = exact (pair, step) tests one window may cost (0 = a window with
# a pair that is not provably clear is blocked, the behaviour before), OSR_SWEEP_KEEP_RMAX = largest swept radius of a
# window that keeps pairs.  Segment parity tests first, then the segment legs per setting ("resolve:rmax" pairs).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "segment or transit or swept or walk or clearance or visibility or touch" > gpurun_out/resolve_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/resolve_pytest.log
for rs in ${RESOLVE_SET:-0:0 512:3e-4}; do
  r=${rs%%:*}; m=${rs##*:}
  echo "== OSR_SWEEP_RESOLVE=$r OSR_SWEEP_KEEP_RMAX=$m"
  export OSR_SWEEP_RESOLVE=$r OSR_SWEEP_KEEP_RMAX=$m
  timeout 600 python scripts/touch_probe.py 2>&1 | tail -1
  timeout 900 python scripts/bench_configs.py --legs ${LEGS:-touch,touch_fine,edges,clearance,ompl_motions} --scale ${SCALE:-1.0} 2>/dev/null | grep "^{" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    keys = [k for k in d if k.endswith('_per_s') or k.startswith('same_') or k.startswith('labels_equal') or k.startswith('labels_and')]
    print(d['leg'], {k: (float('%.4g' % d[k]) if isinstance(d[k], float) else d[k]) for k in keys}, d.get('properties_full_size', {}).get('steps_checksum'), d.get('properties_full_size', {}).get('label_checksum'))
"
done
