#!/usr/bin/env python3
"""Summarise an `ncu --page raw --csv` (and optional `--page source --csv`) export: key metrics, stall mix, opcode mix."""
import csv
import sys
from collections import Counter

raw = list(csv.reader(open(sys.argv[1])))
hdr, units = raw[0], raw[1]
KEYS = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.avg',
        'sm__inst_executed_pipe_tensor.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'sm__cycles_active.avg']
for r in raw[2:]:
    for k in KEYS:
        if k in hdr:
            print(f"{k} = {r[hdr.index(k)]} {units[hdr.index(k)]}")
    st = []
    for i, h in enumerate(hdr):
        if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio'):
            try:
                st.append((float(r[i]), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
            except ValueError:
                pass
    print("stalls/issue:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:9]))
    print('---')
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
    for a, b in zip(starts[:-1], starts[1:]):
        print(rows[a][1][:80])
        h, body = rows[a + 1], rows[a + 2:b]
        ia, isrc, isamp = h.index("Instructions Executed"), h.index("Source"), h.index("# Samples")
        c, s = Counter(), Counter()
        for r in body:
            if len(r) <= ia or not r[ia].isdigit():
                continue
            tok = r[isrc].split()
            op = (tok[1] if tok[0].startswith('@') else tok[0]).split('.')[0]
            c[op] += int(r[ia])
            s[op] += int(r[isamp]) if r[isamp].isdigit() else 0
        tot = sum(c.values())
        print("total warp-instructions", tot)
        print("  ".join(f"{op}:{100 * n / tot:.1f}%/{s[op]}" for op, n in c.most_common(16)))
