for R in 3 4 5 6 8; do echo "== DEC_RAW=$R"; NF4MOE_DEC_RAW=$R timeout 200 python scripts/kernel_times.py --tokens 16 64 --absmax nested --regime gemm_fast 2>&1 | grep -v "^#"; done
