export LIB=build/libnf4moe_exp.so
for cfg in "8 6 2880,0,40" "3 6 2880,0,40" "2 6 2880,0,40" "3 6 2880,1,40" "8 3 2880,1,40" "3 3 2880,1,40" "3 6 2880,1,8" "4 3 3072,1,40"; do
  set -- $cfg
  echo "=== MAXRAW=$1 MAXB=$2 $3"; NF4MOE_DEC_MAXRAW=$1 NF4MOE_DEC_MAXB=$2 timeout 120 python scripts/debug_k4096.py $3 2>&1 | grep -E "FAIL|status|refused"
done
