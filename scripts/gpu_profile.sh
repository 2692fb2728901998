#!/bin/bash
# One gpurun call's worth of evidence: role counters, kernel timings, the ncu launch list of bench.py and one
# `ncu --set full` capture each of the decode GEMV and the prefill tcgen05 GEMM.  Every ncu command runs only after the
# same command line exited 0 without ncu (B200_PROFILING.md).  Outputs land in gpurun_out/.
set -u
O=gpurun_out
mkdir -p $O
TAG=${1:-r1}
python scripts/gemm_roles.py --tokens 8192 > $O/${TAG}_gemm_roles.txt 2>&1
python scripts/kernel_times.py --tokens 1 8 16 32 64 --absmax nested flat > $O/${TAG}_kernel_times.txt 2>&1
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$BENCH > $O/${TAG}_bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv $BENCH > $O/${TAG}_ncu_bench.log 2>&1
DEC="python scripts/prof_decode.py --tokens 64 --layers 2 --iters 2"
$DEC > $O/${TAG}_dec_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 4 -c 2 -f -o $O/${TAG}_gemv $DEC > $O/${TAG}_ncu_gemv.log 2>&1
PRE="python scripts/prof_decode.py --tokens 8192 --layers 1 --iters 2 --regime gemm --absmax flat"
$PRE > $O/${TAG}_pre_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 2 -f -o $O/${TAG}_gemm $PRE > $O/${TAG}_ncu_gemm.log 2>&1
tail -n 12 $O/${TAG}_gemm_roles.txt $O/${TAG}_kernel_times.txt
tail -n 2 $O/${TAG}_dec_plain.log $O/${TAG}_pre_plain.log $O/${TAG}_ncu_gemv.log $O/${TAG}_ncu_gemm.log
ls -la $O
