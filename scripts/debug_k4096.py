#!/usr/bin/env python3
"""Scratch diagnostic: large-K shapes through the decode tcgen05 kernel, synchronising after every launch."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("gpt-oss-vllm_b200")
if os.environ.get("LIB"):
    pkg.load_library(path=os.path.abspath(os.environ["LIB"]))
dev = torch.device("cuda:0")
def step(name, fn):
    try:
        r = fn(); torch.cuda.synchronize(); print("ok  ", name, flush=True); return r
    except Exception as e:
        print("FAIL", name, type(e).__name__, str(e)[:60], flush=True); sys.exit(1)
import subprocess
if len(sys.argv) == 1:
    for c in ["2880,0,40", "4096,1,40", "3072,0,40", "3072,0,8", "3328,0,40", "3584,0,40", "4096,0,8"]:
        subprocess.run([sys.executable, __file__, c])
    sys.exit(0)
HH, nn, TT = map(int, sys.argv[1].split(","))
for (H, nested, T) in [(HH, bool(nn), TT)]:
    E, I, k = 2, 64, 1
    print(f"--- H={H} nested={nested} T={T}", flush=True)
    w = (torch.randn(E, 2 * I, H, device=dev) * 0.02).half()
    from oracle import nf4_oracle as orc
    w13 = pkg.NF4ExpertWeights.quantize(w, torch.from_numpy(orc.create_dynamic_map()))
    ids = torch.zeros(T, 1, dtype=torch.int32, device=dev)
    x = torch.randn(T, H, device=dev).half()
    b = step("bucket", lambda: pkg.ops.bucket(ids, E))
    print("meta", b["meta"][:5].tolist(), flush=True)
    xs = step("permute", lambda: pkg.ops.permute_rows(x, b, k))
    try:
        raw = step("w13", lambda: pkg.ops.expert_w13(xs, w13, b, T, k, act=pkg.ACT_CHUNK_SILU, regime=pkg.REGIME_GEMM_FAST, nested=nested))
    except pkg.Nf4MoeError as e:
        print("refused:", e); continue
    ref = x.float() @ w13.dequantize(nested=nested)[0].float().T
    print("status", b["meta"][4].item(), "maxdiff", (raw[:T].float() - ref).abs().max().item(), flush=True)
