set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r1c_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/r1c_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r1c_bench.json 2> $O/r1c_bench.err; echo "bench rc=$?"; head -c 2500 $O/r1c_bench.json; tail -n 5 $O/r1c_bench.err
timeout 300 python scripts/kernel_times.py --tokens 1 8 16 32 64 128 256 --absmax nested flat --regime gemm_fast > $O/r1c_kernel_times_fast.txt 2>&1
timeout 300 python scripts/kernel_times.py --tokens 1 8 16 32 64 --absmax nested flat --regime gemv > $O/r1c_kernel_times_gemv.txt 2>&1
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-prefill"
$BENCH > $O/r1c_bench_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/r1c_launches.csv $BENCH > $O/r1c_ncu_bench.log 2>&1
DEC="python scripts/prof_decode.py --tokens 64 --layers 1 --iters 2 --regime auto --absmax nested"
$DEC > $O/r1c_dec_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dec_kernel -s 2 -c 2 -f -o $O/r1c_dec $DEC > $O/r1c_ncu_dec.log 2>&1
tail -n 2 $O/r1c_dec_plain.log $O/r1c_ncu_dec.log
cat $O/r1c_kernel_times_fast.txt
