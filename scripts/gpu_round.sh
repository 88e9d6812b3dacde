#!/bin/bash
# GPU-box visit used during development: parity tests, then touch / segment legs for the K2 variants.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
OSR_SEG_WARPS=8 timeout 900 python -m pytest tests -m gpu -x -q -k "segment or edge or visib" 2>&1 | tail -3
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["e2e"]["value"], d["touch"])'
echo "seg 12 warps"; timeout 300 python bench.py --no-cpu-baseline --steps 10 --warmup 3 | python -c "$P"
echo "seg 8 warps"; OSR_SEG_WARPS=8 timeout 300 python bench.py --no-cpu-baseline --steps 10 --warmup 3 | python -c "$P"
echo "tile"; OSR_VALID_KERNEL=tile timeout 300 python bench.py --no-cpu-baseline --steps 10 --warmup 3 | python -c "$P"
timeout 1200 python scripts/bench_configs.py --scale ${SCALE:-0.25} --legs ${CONFIGS:-touch,touch_fine,edges,contacts} > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; cat gpurun_out/configs.jsonl; tail -5 gpurun_out/configs.err
