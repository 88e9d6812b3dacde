#!/bin/bash
# GPU-box visit used during development: parity tests, e2e chunk sweep, bounded configs 2-5.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for c in 32768 65536 131072 262144; do
  echo "feed chunk $c"; OSR_HOST_CHUNK=$c timeout 300 python bench.py --no-touch --no-cpu-baseline --steps 20 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
done
echo "no feed"; OSR_NO_STREAM_FEED=1 OSR_HOST_CHUNK=524288 timeout 300 python bench.py --no-touch --no-cpu-baseline --steps 20 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
if [[ -n "$CONFIGS" ]]; then
timeout 1200 python scripts/bench_configs.py --scale ${SCALE:-0.25} --legs $CONFIGS > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; cat gpurun_out/configs.jsonl; tail -5 gpurun_out/configs.err
fi
