#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2c}
timeout 200 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; echo "debug rc=$?"; cat $O/${TAG}_debug.txt | tail -n 30
timeout 1500 python -m pytest tests -m gpu -q --timeout=300 --timeout-method=thread --durations=10 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" $O/${TAG}_pytest.log | head -60
