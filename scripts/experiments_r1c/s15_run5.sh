set -u
for L in libnf4moe_old.so libnf4moe_nopin.so libnf4moe.so; do
  echo "== $L"
  NF4MOE_LIB=$PWD/gpt-oss-vllm_b200/$L timeout 200 python scripts/kernel_times.py --tokens 16 32 64 --absmax nested --regime gemm_fast 2>&1 | tail -n 6
done
echo "== libnf4moe.so sleep 0"
NF4MOE_DEC_SLEEP=0,0,0,0 timeout 200 python scripts/kernel_times.py --tokens 16 32 64 --absmax nested --regime gemm_fast 2>&1 | tail -n 6
echo "== libnf4moe.so again"
timeout 200 python scripts/kernel_times.py --tokens 16 32 64 --absmax nested --regime gemm_fast 2>&1 | tail -n 6
