#!/bin/bash
# Lean A/B run for decode-kernel changes: small-shape check, layer times, role counters, the decode parity tests.
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-exp}
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; rc=$?; echo "debug rc=$rc $(grep BAD $O/${TAG}_debug.txt | head -3)"
if [ $rc -ne 0 ]; then tail -5 $O/${TAG}_debug.txt; echo "stopping"; exit 1; fi
timeout 300 python scripts/layer_times.py --tokens 16 64 128 256 > $O/${TAG}_lt.txt 2>&1; echo "lt rc=$?"; cat $O/${TAG}_lt.txt
timeout 200 python scripts/layer_times.py --tokens 64 --regime gemm > $O/${TAG}_lt_exact.txt 2>&1; cat $O/${TAG}_lt_exact.txt
timeout 200 python scripts/dec_roles.py --tokens 64 > $O/${TAG}_dec_roles.txt 2>&1; tail -n 8 $O/${TAG}_dec_roles.txt
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -x -q --timeout=300 --timeout-method=thread > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/${TAG}_pytest.log
