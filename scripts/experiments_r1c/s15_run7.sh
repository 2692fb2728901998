set -u
O=gpurun_out; mkdir -p $O
export NF4MOE_LIB=$PWD/gpt-oss-vllm_b200/libnf4moe_g3.so
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "gemm" > $O/s15_pytest7.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/s15_pytest7.log
echo "== G3"
timeout 200 python scripts/kernel_times.py --tokens 1 8 16 32 64 128 --absmax nested --regime gemm_fast 2>&1 | tail -n 8
unset NF4MOE_LIB
echo "== G2"
timeout 200 python scripts/kernel_times.py --tokens 1 8 16 32 64 128 --absmax nested --regime gemm_fast 2>&1 | tail -n 8
