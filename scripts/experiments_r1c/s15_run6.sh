set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "gemm" > $O/s15_pytest6.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/s15_pytest6.log
for B in 3 4 5 6; do
  echo "== DEC_B=$B"
  NF4MOE_DEC_B=$B timeout 200 python scripts/kernel_times.py --tokens 1 8 16 32 64 128 --absmax nested --regime gemm_fast 2>&1 | tail -n 4
done
