set -u
O=gpurun_out; mkdir -p $O
DEC="python scripts/prof_decode.py --tokens 256 --layers 1 --iters 2 --regime auto --absmax nested"
$DEC > $O/r2p_dec256_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:dec_kernel -s 1 -c 1 -f -o $O/r2p_dec256 $DEC > $O/r2p_ncu_dec256.log 2>&1
tail -n 1 $O/r2p_dec256_plain.log $O/r2p_ncu_dec256.log
G1="python scripts/prof_decode.py --tokens 1 --layers 1 --iters 2 --regime auto --absmax nested"
$G1 > $O/r2p_t1_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/r2p_t1_launches.csv $G1 > $O/r2p_ncu_t1.log 2>&1
tail -n 1 $O/r2p_t1_plain.log
