#!/bin/bash
# ncu --set full of the prefill tcgen05 grouped GEMM (T=8192, EXACT and FAST operands) and of the decode GEMV (T=8); each ncu run
# follows a plain run of the same command line that exited 0.  Usage: bash scripts/gpu_profile_prefill.sh [tag]
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r1c}
for MODE in gemm gemm_fast; do
PRE="python scripts/prof_decode.py --tokens 8192 --layers 1 --iters 2 --regime $MODE --absmax nested"
$PRE > $O/${TAG}_pre_${MODE}_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 2 -f -o $O/${TAG}_prefill_$MODE $PRE > $O/${TAG}_ncu_pre_$MODE.log 2>&1
tail -n 1 $O/${TAG}_pre_${MODE}_plain.log $O/${TAG}_ncu_pre_$MODE.log
done
GV="python scripts/prof_decode.py --tokens 8 --layers 1 --iters 2 --regime auto --absmax nested"
$GV > $O/${TAG}_gemv_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 2 -c 2 -f -o $O/${TAG}_gemv_T8 $GV > $O/${TAG}_ncu_gemv.log 2>&1
tail -n 1 $O/${TAG}_gemv_plain.log $O/${TAG}_ncu_gemv.log
timeout 300 python scripts/prof_prefill.py --tokens 8192 --absmax nested --regime gemm > $O/${TAG}_prefill_times.txt 2>&1
timeout 300 python scripts/prof_prefill.py --tokens 8192 --absmax nested --regime gemm_fast >> $O/${TAG}_prefill_times.txt 2>&1
cat $O/${TAG}_prefill_times.txt | tail -n 4
