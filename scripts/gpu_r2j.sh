#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2j}
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; rc=$?; echo "debug rc=$rc $(grep BAD $O/${TAG}_debug.txt)"
if [ $rc -ne 0 ]; then tail -5 $O/${TAG}_debug.txt; echo "stopping"; exit 1; fi
LT="python scripts/layer_times.py"
timeout 200 $LT --tokens 16 64 128 256 > $O/${TAG}_lt.txt 2>&1; echo "lt rc=$?"; cat $O/${TAG}_lt.txt
NF4MOE_DEC_MERGED=0 timeout 200 $LT --tokens 16 64 128 256 --lib build/libnf4moe_exp.so --tag "split " > $O/${TAG}_lt_split.txt 2>&1; cat $O/${TAG}_lt_split.txt
timeout 200 $LT --tokens 64 --regime gemm > $O/${TAG}_lt_exact.txt 2>&1; cat $O/${TAG}_lt_exact.txt
timeout 1300 python -m pytest tests -m gpu -x -q --timeout=400 --timeout-method=thread > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/${TAG}_pytest.log
