#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2e}
timeout 200 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; echo "debug rc=$? $(grep BAD $O/${TAG}_debug.txt)"
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_round2.py -q -x -k "cannot_place" > $O/${TAG}_sanitizer.txt 2>&1; echo "sanitizer rc=$?"; grep -E "Invalid|Error|error|at 0x|by thread|Address" $O/${TAG}_sanitizer.txt | head -30
timeout 1500 python -m pytest tests -m gpu -q --timeout=300 --timeout-method=thread --durations=10 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" $O/${TAG}_pytest.log | head -60
