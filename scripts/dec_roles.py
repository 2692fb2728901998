#!/usr/bin/env python3
"""Where the warp roles of the decode tcgen05 kernel (gemm_dec.cuh) spend their cycles: CTA 0, clock64 counters, w13."""
import argparse, ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc
p = argparse.ArgumentParser(); p.add_argument("--tokens", type=int, nargs="+", default=[64])
p.add_argument("--absmax", default="nested"); a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
lib = pkg.load_library(path=os.path.join(ROOT, "build", "libnf4moe_exp.so"))   # python gpt-oss-vllm_b200/build.py -DNF4MOE_EXPERIMENTS --out=build/libnf4moe_exp.so
dev = torch.device("cuda:0"); H = I = 2880; E, k = 32, 4
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
dbg = torch.zeros(64, dtype=torch.int64, device=dev)
names = {0: ("dequant w0", ["wait m_full", "wait w_full", "decode kb0", "wait a_empty", "st+decode kb1+st", "wait::st+arrive"]),
         8: ("dequant w9", ["wait m_full", "wait w_full", "decode kb0", "wait a_empty", "st+decode kb1+st", "wait::st+arrive"]),
         16: ("MMA", ["wait d_empty", "wait w_full", "wait a_full", "issue+commit"]),
         24: ("producer", ["absmax", "wait w_empty", "issue"])}
NESTED = a.absmax == "nested"
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half(); lg = torch.randn(T, E, generator=g, device=dev).half()
    ids, wts = pkg.ops.route_topk(lg, k); b = pkg.ops.bucket(ids, E); xs = pkg.ops.permute_rows(x, b, k)
    for _ in range(3):
        pkg.ops.expert_w13(xs, w13, b, T, k, regime=pkg.REGIME_GEMM_FAST, nested=NESTED)
    torch.cuda.synchronize()
    lib.nf4moe_debug_set_gemm_counters(ctypes.c_void_p(dbg.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); pkg.ops.expert_w13(xs, w13, b, T, k, regime=pkg.REGIME_GEMM_FAST, nested=NESTED); e1.record(); torch.cuda.synchronize()
    lib.nf4moe_debug_set_gemm_counters(ctypes.c_void_p(0))
    d = dbg.cpu().tolist()
    ntile = int(b["meta"][2].item()); items0 = (ntile * 45 + 147) // 148
    nstage = items0 * 12
    print(f"T={T} {a.absmax}: w13 {e0.elapsed_time(e1) * 1e3:.1f} us, token tiles {ntile}, ~{items0} items on CTA 0 ({nstage} stages)")
    for nm, off in (("CTA 0", 32), ("last CTA", 48)):
        t = d[off:off + 10]
        lab = ["entry", "prologue done", "dep wait done", "setup done", "first raw stage landed", "first A operand ready",
               "first accumulator done", "last accumulator done", "epilogue stores done", "exit"]
        print(f"  timeline {nm} (us after CTA 0 entered): " + ", ".join(f"{l} {((v - d[32]) / 1e3):.2f}" for l, v in zip(lab, t) if v))
    for base, (role, labels) in names.items():
        tot = sum(d[base:base + 6])
        print(f"  {role:11s} " + "  ".join(f"{l}={d[base + i] / nstage:7.1f}" for i, l in enumerate(labels)) + f"   (cycles per CTA stage; sum {tot / nstage:.1f})")
