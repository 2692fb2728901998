#!/usr/bin/env python3
"""SASS listings of the hot kernels of the shipped libnf4moe.so -> profiles/r3_sass/*.sass (+ README with mnemonic counts),
and profiles/sass_mnemonics.txt: the same counts for EVERY kernel of the library (whole functions, no line cap).
`cuobjdump -sass`, encoding columns stripped, one file per kernel."""
import os, re, subprocess, sys
from collections import Counter
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gpt-oss-vllm_b200", "libnf4moe.so")
OUT = os.path.join(ROOT, "profiles", "r3_sass")
WANT = {  # mangled-name fragment -> file name
    "10dec_kernelILi2ELb1ELb0": "dec_kernel_w13_w2_fast.sass", "10dec_kernelILi2ELb0ELb0": "dec_kernel_w13_w2_exact.sass",
    "10dec_kernelILi0ELb1ELb0": "dec_kernel_w13_fast.sass", "10dec_kernelILi1ELb1ELb0": "dec_kernel_w2_fast.sass",
    "11gemm_kernelILb1ELb0": "gemm_kernel_w13_exact.sass", "11gemm_kernelILb1ELb1": "gemm_kernel_w13_fast.sass",
    "11gemv_kernelILi1ELi9ELb1ELb1ELi45": "gemv_kernel_w13_nested_k2880.sass", "13bucket_kernel": "bucket_kernel.sass",
    "17route_topk_kernelILi1": "route_topk_kernel.sass", "14combine_kernel": "combine_kernel.sass",
    "18ep_dispatch_kernel": "ep_dispatch_kernel.sass", "16ep_return_kernel": "ep_return_kernel.sass",
    "25route_bucket_small_kernelILi1": "route_bucket_small_kernel.sass", "21bucket_cluster_kernel": "bucket_cluster_kernel.sass",
    "19permute_rows_kernel": "permute_rows_kernel.sass",
}
KEYS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UBLKCP", "UTMAPF", "STTM", "LDTM", "SYNCS", "LDS", "PRMT", "HMUL2", "HFMA2", "FMUL", "F2FP", "LDG",
        "STG", "STL", "LDL", "USETMAXREG", "HMMA", "SHFL", "REDG"]
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
os.makedirs(OUT, exist_ok=True)
for f in os.listdir(OUT):
    if f.endswith(".sass"):
        os.remove(os.path.join(OUT, f))
readme = ["# SASS listings of the shipped libnf4moe.so (round 3 build)", "",
          "`python scripts/make_sass_listings.py`: `cuobjdump -sass gpt-oss-vllm_b200/libnf4moe.so` with the encoding columns stripped, one file per "
          "kernel named below (static instruction counts; UTCHMMA = tcgen05.mma, STTM/LDTM = tcgen05.st/ld, UTMALDG = TMA tensor load, UBLKCP = TMA "
          "bulk copy, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, USETMAXREG = setmaxnreg, STL/LDL = local-memory traffic, HMMA = mma.sync).", ""]
allk = ["# SASS mnemonic counts of EVERY kernel of libnf4moe.so (scripts/make_sass_listings.py: cuobjdump -sass, whole functions).",
        "# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG / UBLKCP / UTMAPF = TMA tensor load / bulk copy / prefetch, UTCBAR = tcgen05.commit,",
        "# HMMA = mma.sync, SYNCS = mbarrier ops, STL/LDL = local-memory traffic."]
for fn in txt.split("Function : ")[1:]:
    name = fn.split("\n", 1)[0].strip()
    hit = [v for k, v in WANT.items() if k in name]
    lines, ops = [], Counter()
    for l in fn.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", l)
        if m:
            lines.append(f"/*{m.group(1)}*/  {m.group(2).rstrip()}")
            mm = re.match(r"(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", m.group(2).strip())
            if mm:
                ops[mm.group(1)] += 1
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0]
    allk.append(f"{dem}\n    total {len(lines)}: " + " ".join(f"{k}={ops[k]}" for k in KEYS if ops[k]))
    if not hit:
        continue
    open(os.path.join(OUT, hit[0]), "w").write(f"// {dem}\n" + "\n".join(lines) + "\n")
    readme.append(f"* `{hit[0]}` -- `{dem}`: {len(lines)} instructions; " + ", ".join(f"{k} {ops[k]}" for k in KEYS if ops[k]))
open(os.path.join(OUT, "README.md"), "w").write("\n".join(readme) + "\n")
open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w").write("\n".join(allk) + "\n")
print("\n".join(readme[3:]))
