set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "gemm" > $O/s15_pytest3.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/s15_pytest3.log
for S in "0,0,0,0" "500,100,40,0" "500,100,0,0" "1000,200,40,40" "500,100,40,100" "2000,300,100,0"; do
  echo "SLEEP=$S"
  NF4MOE_DEC_SLEEP=$S timeout 200 python scripts/kernel_times.py --tokens 16 64 --absmax nested --regime gemm_fast 2>&1 | tail -n 2
done
