set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/s15_pytest.log 2>&1; echo "pytest rc=$?" 
tail -5 $O/s15_pytest.log
timeout 300 python scripts/kernel_times.py --tokens 1 8 16 32 64 --absmax nested --regime gemv > $O/s15_kt_gemv.txt 2>&1
timeout 300 python scripts/kernel_times.py --tokens 1 8 16 32 64 128 --absmax nested --regime gemm_fast > $O/s15_kt_fast.txt 2>&1
tail -8 $O/s15_kt_gemv.txt $O/s15_kt_fast.txt
timeout 600 python bench.py --steps 10 --warmup 3 > $O/s15_bench.json 2> $O/s15_bench.err; echo "bench rc=$?"
tail -c 1500 $O/s15_bench.json
