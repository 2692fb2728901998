#!/usr/bin/env python3
"""Turn what gpurun brought back (gpurun_out/*.ncu-rep, *_launches.csv) into the small tracked summaries under profiles/,
and list the SASS mnemonics of every kernel in libnf4moe.so (the proof of tcgen05 / TMA use, B200_PROFILING.md).

    python scripts/make_profiles.py <tag> [<rep-basename> <out-name> <note>]...
    python scripts/make_profiles.py --sass
"""
import csv
import json
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
O, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
KEYS = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.avg', 'sm__cycles_active.avg',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'launch__grid_size', 'launch__block_size',
        'launch__shared_mem_per_block_dynamic']


def summarise(rep: str, out: str, note: str) -> None:
    raw_csv = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    raw = list(csv.reader(l for l in raw_csv.splitlines() if l.startswith('"')))
    h, u = raw[0], raw[1]
    res = []
    for r in raw[2:]:
        d = {k: (r[h.index(k)] + ' ' + u[h.index(k)]).strip() for k in KEYS if k in h}
        st = []
        for i, n in enumerate(h):
            if n.startswith('smsp__average_warps_issue_stalled') and n.endswith('per_issue_active.ratio'):
                try:
                    st.append((float(r[i]), n.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
                except ValueError:
                    pass
        d['stalls_per_issue'] = {n: round(v, 2) for v, n in sorted(st, reverse=True)[:8]}
        res.append(d)
    json.dump({'note': note, 'kernels': res}, open(out, 'w'), indent=1)
    print("wrote", out)


def sass() -> None:
    lib = os.path.join(ROOT, "gpt-oss-vllm_b200", "libnf4moe.so")
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    out = ["# SASS mnemonic counts per kernel of libnf4moe.so (cuobjdump -sass).  UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st,",
           "# UTMALDG / UBLKCP / UTMAPF = TMA tensor load / bulk copy / tensor prefetch, HMMA = mma.sync, SYNCS = mbarrier."]
    for fn in txt.split("Function : ")[1:]:
        name = fn.split("\n", 1)[0].strip()
        ops = Counter()
        for l in fn.split("\n"):
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", l)
            if m:
                ops[m.group(1)] += 1
        demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
        keys = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UBLKCP", "HMMA", "SYNCS", "LDS", "PRMT", "HMUL2", "HFMA2",
                "FMUL", "F2FP", "LDG", "STG", "SHFL", "MUFU"]
        out.append(f"{demangled[:110]}\n    total {sum(ops.values())}: " + " ".join(f"{k}={ops[k]}" for k in keys if ops[k]))
    open(os.path.join(P, "sass_mnemonics.txt"), "w").write("\n".join(out) + "\n")
    print("wrote profiles/sass_mnemonics.txt")


if __name__ == "__main__":
    os.makedirs(P, exist_ok=True)
    if "--sass" in sys.argv:
        sass()
    else:
        args = sys.argv[1:]
        for i in range(0, len(args) - 2, 3):
            summarise(os.path.join(O, args[i]), os.path.join(P, args[i + 1]), args[i + 2])
