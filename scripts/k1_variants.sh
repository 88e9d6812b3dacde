#!/bin/bash
# A/B runs of K1 builds (puzzletunneldiscovery_b200/build.py --variant TAG -D...): one bench line per library.
#   usage: scripts/k1_variants.sh TAG [TAG ...]     ("base" = the product library)
mkdir -p gpurun_out
for tag in "$@"; do
  lib=puzzletunneldiscovery_b200/libosrgpu_$tag.so
  [[ $tag == base ]] && lib=puzzletunneldiscovery_b200/libosrgpu.so
  OSR_LIB=$PWD/$lib timeout 300 python bench.py --no-cpu-baseline --no-touch --steps 40 --warmup 5 > gpurun_out/var_$tag.json 2> gpurun_out/var_$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/var_%s.json" % tag).read().strip().splitlines()[-1])
    c = d["engine_counters_per_state"]
    print("%-10s value %.4e  kernel_ms %.4f  e2e %.3e  e2e32 %.3e  bv %.2f leaf %.3f fp64 %.3f" % (
        tag, d["value"], d["roofline"]["kernel_ms"], d["e2e"]["value"], d["e2e_f32_states"]["value"],
        c["bv_tests"], c["leaf_tests"], c["fp64_rechecks"]))
except Exception as e:
    print(tag, "FAILED", e)
PY
done
