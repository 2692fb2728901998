LT="python scripts/layer_times.py --lib build/libnf4moe_exp.so --tokens 64 --absmax nested flat"
timeout 200 $LT --tag "rowmajor "
NF4MOE_DEC_FAKE_TILED=1 timeout 200 $LT --tag "contig16K "
