#!/usr/bin/env python3
"""Where each warp role of the tcgen05 kernel spends its cycles (CTA 0, clock64 counters): w13 GEMM only."""
import argparse, ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc
p = argparse.ArgumentParser(); p.add_argument("--tokens", type=int, nargs="+", default=[8192])
p.add_argument("--regime", default="gemm", choices=["gemm", "gemm_fast"]); p.add_argument("--absmax", default="flat"); a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
lib = pkg.load_library()
dev = torch.device("cuda:0"); H = I = 2880; E, k = 32, 4
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
dbg = torch.zeros(64, dtype=torch.int64, device=dev)
names = {0: ("B-producer", ["wait b_empty", "issue"]), 8: ("W-producer", ["wait w_empty"]),
         16: ("MMA", ["wait d_empty", "wait full", "-", "issue+commit+next addr"]),
         24: ("dequant g0", ["wait m_full", "wait w_full", "decode", "wait a_empty", "st+arrive"]),
         32: ("dequant g1", ["wait m_full", "wait w_full", "decode", "wait a_empty", "st+arrive"]),
         40: ("dequant g2", ["wait m_full", "wait w_full", "decode", "wait a_empty", "st+arrive"]),
         56: ("epilogue", ["wait d_full", "ld+act+store"])}
REG = pkg.REGIME_GEMM_FAST if a.regime == "gemm_fast" else pkg.REGIME_GEMM
NESTED = a.absmax == "nested"
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half(); lg = torch.randn(T, E, generator=g, device=dev).half()
    ids, wts = pkg.ops.route_topk(lg, k); b = pkg.ops.bucket(ids, E); xs = pkg.ops.permute_rows(x, b, k)
    pkg.ops.expert_w13(xs, w13, b, T, k, regime=REG, nested=NESTED); torch.cuda.synchronize()
    lib.nf4moe_debug_set_gemm_counters(ctypes.c_void_p(dbg.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); pkg.ops.expert_w13(xs, w13, b, T, k, regime=REG, nested=NESTED); e1.record(); torch.cuda.synchronize()
    lib.nf4moe_debug_set_gemm_counters(ctypes.c_void_p(0))
    d = dbg.cpu().tolist()
    ntile = int(b["meta"][2].item()); items0 = (ntile * 45 + 147) // 148
    nrt = 45
    print(f"T={T} {a.regime} {a.absmax}: w13 {e0.elapsed_time(e1) * 1e3:.1f} us, token tiles {ntile}, ~{items0} items on CTA 0 ({items0 * 45} k-blocks)")
    for base, (role, labels) in names.items():
        tot = sum(d[base:base + 6])
        print(f"  {role:11s} " + "  ".join(f"{l}={d[base + i] / max(items0 * 45, 1):8.1f}" for i, l in enumerate(labels)) + f"   (cycles per k-block; sum {tot / max(items0 * 45, 1):.1f})")
