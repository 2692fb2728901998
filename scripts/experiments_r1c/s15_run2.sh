set -u
O=gpurun_out; mkdir -p $O
DEC="python scripts/prof_decode.py --tokens 64 --layers 1 --iters 2 --regime gemm_fast --absmax nested"
$DEC > $O/s15_dec_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dec_kernel -s 2 -c 2 -f -o $O/s15_dec $DEC > $O/s15_ncu_dec.log 2>&1
timeout 120 python scripts/dec_roles.py > $O/s15_dec_roles.txt 2>&1
tail -3 $O/s15_dec_plain.log $O/s15_ncu_dec.log; tail -40 $O/s15_dec_roles.txt
