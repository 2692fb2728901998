#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-exp5}
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; rc=$?; echo "debug rc=$rc $(grep BAD $O/${TAG}_debug.txt | head -3)"
if [ $rc -ne 0 ]; then tail -8 $O/${TAG}_debug.txt; fi
grep "E=4 H=192 I=128 T=40" $O/${TAG}_debug.txt | head -4
LT="python scripts/layer_times.py --tokens 128 192 256 320 --reps 6 --absmax nested"
timeout 300 $LT --tag "auto-tiles " > $O/${TAG}_lt.txt 2>&1; cat $O/${TAG}_lt.txt
NF4MOE_TILE=32 timeout 300 $LT --lib build/libnf4moe_exp.so --tag "32-row-tiles " > $O/${TAG}_lt32.txt 2>&1; cat $O/${TAG}_lt32.txt
NF4MOE_TILE=64 timeout 300 python scripts/layer_times.py --tokens 128 384 512 --reps 6 --absmax nested --regime gemm_fast --lib build/libnf4moe_exp.so --tag "64-row-tiles " > $O/${TAG}_lt64.txt 2>&1; cat $O/${TAG}_lt64.txt
NF4MOE_TILE=160 timeout 300 python scripts/layer_times.py --tokens 256 384 512 --reps 6 --absmax nested --regime gemm_fast --lib build/libnf4moe_exp.so --tag "160-row-tiles " > $O/${TAG}_lt160.txt 2>&1; cat $O/${TAG}_lt160.txt
timeout 900 python -m pytest tests -m gpu -x -q --timeout=300 --timeout-method=thread > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/${TAG}_pytest.log
