#!/bin/bash
# GPU-box visit: full check (tests, smoke, bench both arms, ncu launch list + full capture of K1) and the bounded
# runs of configs 2-5 at full scale.
bash scripts/gpu_check.sh all
timeout 1500 python scripts/bench_configs.py --scale 1.0 > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; cut -c1-600 gpurun_out/configs.jsonl; tail -3 gpurun_out/configs.err
