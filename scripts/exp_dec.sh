#!/bin/bash
# sweep of the decode tcgen05 kernel's ring depths
for cfg in "8 4" "4 4" "8 2" "3 3" "8 6"; do
  set -- $cfg
  echo "=== RAW=$1 B=$2 ==="
  NF4MOE_DEC_RAW=$1 NF4MOE_DEC_B=$2 timeout 120 python scripts/kernel_times.py --regime gemm_fast --tokens 16 64 --absmax nested flat 2>&1 | tail -4
done
