#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, then (optionally) the ncu launch list and a full capture
# of the validity kernel.  Everything is written under gpurun_out/.
#   usage: scripts/gpu_check.sh [tests|bench|ncu|all]
what=${1:-all}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
if [[ $what == all || $what == tests ]]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
  tail -15 gpurun_out/pytest_gpu.log
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
fi
if [[ $what == all || $what == bench ]]; then
  timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
  timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
fi
if [[ $what == all || $what == ncu ]]; then
  CMD="python bench.py --steps 3 --warmup 1 --no-cpu-baseline --touch-segments 4096"
  timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
  CMD2="python bench.py --steps 3 --warmup 1 --no-cpu-baseline --no-touch"
  timeout 600 $CMD2 > gpurun_out/ncu_plain2.log 2>&1 &&
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_valid_dfs -s 2 -c 1 -f -o gpurun_out/prof_valid $CMD2 > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
fi
if [[ $what == all || $what == ncu2 ]]; then
  # full capture of the segment kernel on the config-2 touch leg (swept windows on)
  CMD3="python scripts/bench_configs.py --legs touch --scale 0.25"
  timeout 600 $CMD3 > gpurun_out/ncu_plain3.log 2>&1 &&
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_transit -s 1 -c 1 -f -o gpurun_out/prof_transit $CMD3 > gpurun_out/ncu_full_transit.log 2>&1
  echo "ncu transit rc=$?"; tail -3 gpurun_out/ncu_full_transit.log
fi
if [[ $what == all || $what == configs ]]; then
  timeout 1200 python scripts/bench_configs.py 2> gpurun_out/configs.err | grep "^{" > gpurun_out/configs_2_to_5.jsonl; echo "configs lines: $(wc -l < gpurun_out/configs_2_to_5.jsonl)"
fi
