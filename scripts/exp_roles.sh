#!/bin/bash
# A/B of the control-warp placement in the tcgen05 kernel (NF4MOE_GEMM_EXP=16: old placement, warps 0-3).
set -u
O=gpurun_out; mkdir -p $O
for X in 16 0; do
  echo "=== NF4MOE_GEMM_EXP=$X ==="
  NF4MOE_GEMM_EXP=$X python scripts/gemm_roles.py --tokens 8192 2>&1 | tail -9
  NF4MOE_GEMM_EXP=$X python scripts/gemm_roles.py --tokens 64 --regime gemm_fast --absmax nested 2>&1 | tail -9
  NF4MOE_GEMM_EXP=$X python scripts/prof_prefill.py --tokens 8192 --regime gemm 2>&1 | tail -1
  NF4MOE_GEMM_EXP=$X python scripts/prof_prefill.py --tokens 8192 --regime gemm_fast --check 2>&1 | tail -2
  NF4MOE_GEMM_EXP=$X python scripts/kernel_times.py --regime gemm_fast --tokens 16 64 --absmax nested flat 2>&1 | tail -4
done
