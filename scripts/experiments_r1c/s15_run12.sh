for TILE in 32 160; do echo "== TILE=$TILE"; NF4MOE_TILE=$TILE timeout 300 python scripts/kernel_times.py --tokens 128 256 384 512 768 1024 --absmax nested --regime gemm_fast 2>&1 | grep -v "^#"; done
echo "== E=8 (EP4 rank shape), rows as tokens k=4"
for TILE in 32 160; do echo "== TILE=$TILE"; NF4MOE_TILE=$TILE timeout 300 python scripts/kernel_times.py --experts 8 --tokens 64 128 --absmax nested --regime gemm_fast 2>&1 | grep -v "^#"; done
for TILE in 32 160; do echo "== TILE=$TILE E=4"; NF4MOE_TILE=$TILE timeout 300 python scripts/kernel_times.py --experts 4 --tokens 64 128 --absmax nested --regime gemm_fast 2>&1 | grep -v "^#"; done
