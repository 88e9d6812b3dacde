# small-call latency, same box: default against OSR_SMALL_BOTH=0
set -e
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "small or edge_cases" 2>&1 | tail -3
for b in 0 4; do echo "== OSR_SMALL_BOTH=$b"; OSR_SMALL_BOTH=$b python scripts/latency_probe.py 2>&1 | grep -E "n=     1:|n=    32|n=  1024|n=  4096|n= 16384|one (random|free)|is_valid_state "; done
