for cfg in "160 4" "128 4" "128 6" "144 4" "96 8" "112 6"; do set -- $cfg
echo "== TILE=$1 B=$2"
NF4MOE_TILE=$1 NF4MOE_GEMM_B=$2 timeout 200 python scripts/prof_prefill.py --tokens 8192 --absmax nested --regime gemm_fast 2>&1 | tail -n 4
done
