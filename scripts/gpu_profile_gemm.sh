#!/bin/bash
# ncu --set full of the tcgen05 kernel in both regimes (prefill T=8192, decode T=64); each ncu run follows a plain run of
# the same command line that exited 0.
set -u
O=gpurun_out
mkdir -p $O
TAG=${1:-r1b}
MODE=${2:-gemm_fast}
PRE="python scripts/prof_decode.py --tokens 8192 --layers 1 --iters 2 --regime $MODE --absmax nested"
$PRE > $O/${TAG}_pre_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -f -o $O/${TAG}_gemm_prefill $PRE > $O/${TAG}_ncu_pre.log 2>&1
DEC="python scripts/prof_decode.py --tokens 64 --layers 1 --iters 2 --regime $MODE --absmax nested"
$DEC > $O/${TAG}_dec_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -f -o $O/${TAG}_gemm_decode $DEC > $O/${TAG}_ncu_dec.log 2>&1
tail -n 2 $O/${TAG}_pre_plain.log $O/${TAG}_dec_plain.log $O/${TAG}_ncu_pre.log $O/${TAG}_ncu_dec.log
