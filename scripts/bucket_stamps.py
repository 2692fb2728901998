#!/usr/bin/env python3
"""Timeline of the single-CTA bucket kernel inside the route -> bucket chain (experiments build, %globaltimer stamps)."""
import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("gpt-oss-vllm_b200")
lib = pkg.load_library(path=os.path.join(ROOT, "build", "libnf4moe_exp.so"))
dev = torch.device("cuda:0"); E, k = 32, 4
g = torch.Generator(device=dev).manual_seed(1)
names = ["entry->dep wait done", "ids staged", "counted", "scan+lists", "placed"]
for T in (1, 64, 256):
    lg = torch.randn(T, E, generator=g, device=dev).half()
    rows = []
    for rep in range(6):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(); ids, w = pkg.ops.route_topk(lg, k); e1.record(); b = pkg.ops.bucket(ids, E); e2.record()
        torch.cuda.synchronize()
        st = (ctypes.c_longlong * 8)()
        assert lib.nf4moe_debug_bucket_stamps(st) == 0
        rows.append(([ (st[i + 1] - st[i]) / 1e3 for i in range(5)], e1.elapsed_time(e2) * 1e3))
    d, ev = rows[-1]
    print(f"T={T:4d}: " + " | ".join(f"{n} {v:5.2f}" for n, v in zip(names, d)) + f" | total in-kernel {sum(d):5.2f} us, events bucket {ev:5.2f} us", flush=True)
