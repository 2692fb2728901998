#!/usr/bin/env python3
"""Generate golden vectors by RUNNING THE REFERENCE'S OWN PYTHON (read-only at /root/reference) in the build
container.  The reference imports ``bitsandbytes`` at module top (P/streaming_unsloth_bnb.py:8,
P/unsloth_direct_loader.py:12-13, P/unsloth_to_vllm_converter.py:14-16) but the functions exercised here never
call into it, so an empty stub module is enough to import them.  Nothing here runs on the GPU box; the outputs
(``reference_golden.npz`` + ``reference_golden.json``) are committed and read by tests/test_oracle.py.

Usage (build container only):  python tests/golden/make_golden.py
"""
from __future__ import annotations

import hashlib
import importlib.util
import json
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/patches"
sys.path.insert(0, ROOT)

from oracle import nf4_oracle as orc  # noqa: E402


def _stub_bnb():
    bnb = types.ModuleType("bitsandbytes")
    nn = types.ModuleType("bitsandbytes.nn")
    fn = types.ModuleType("bitsandbytes.functional")

    class _Absent:  # any use of real bnb functionality must fail loudly
        def __init__(self, *a, **k):
            raise RuntimeError("bitsandbytes is stubbed: the golden generator must not reach bnb code")

    nn.Linear4bit = nn.Params4bit = nn.Int4Params = _Absent
    fn.QuantState = _Absent
    fn.dequantize_4bit = fn.dequantize_blockwise = _Absent
    bnb.nn, bnb.functional = nn, fn
    sys.modules.update({"bitsandbytes": bnb, "bitsandbytes.nn": nn, "bitsandbytes.functional": fn})
    # vllm.logger is only used for init_logger(); avoid importing all of vLLM.
    if "vllm" not in sys.modules:
        import logging
        vllm = types.ModuleType("vllm")
        vlog = types.ModuleType("vllm.logger")
        vlog.init_logger = logging.getLogger
        vllm.logger = vlog
        sys.modules.update({"vllm": vllm, "vllm.logger": vlog})


def _load(name):
    spec = importlib.util.spec_from_file_location("ref_" + name, os.path.join(REF, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def synth_state(seed: int, n_weights: int, scalar_nested: bool):
    """Seeded synthetic Unsloth-format tensors (P/unsloth_bnb_loader.py:202-222 key set).  tests regenerate these."""
    rng = np.random.default_rng(seed)
    nb = n_weights // 64
    n2 = 1 if scalar_nested else -(-nb // 256)
    return {
        "packed": rng.integers(0, 256, size=n_weights // 2, dtype=np.uint8),
        "absmax_u8": rng.integers(0, 256, size=nb, dtype=np.uint8),
        "nested_absmax": (rng.random(n2, dtype=np.float32) * np.float32(0.05) + np.float32(0.01)),
        "nested_quant_map": orc.create_dynamic_map(),
        "quant_map": orc.NF4_CODE.copy(),
    }


def write_ckpt(path, st, layer, expert, proj):
    import safetensors.torch
    key = f"model.layers.{layer}.mlp.experts.{proj}.{expert}.weight"
    safetensors.torch.save_file({
        key: torch.from_numpy(st["packed"]).reshape(-1, 1),
        key + ".absmax": torch.from_numpy(st["absmax_u8"]),
        key + ".nested_absmax": torch.from_numpy(st["nested_absmax"]),
        key + ".nested_quant_map": torch.from_numpy(st["nested_quant_map"]),
        key + ".quant_map": torch.from_numpy(st["quant_map"]),
    }, path)


def digest(a16: np.ndarray) -> dict:
    flat = np.ascontiguousarray(a16).reshape(-1)
    pos = np.linspace(0, flat.size - 1, 64).astype(np.int64)
    return {"sha256": hashlib.sha256(flat.tobytes()).hexdigest(), "shape": list(a16.shape),
            "sample_pos": pos.tolist(), "sample_bits": flat.view(np.uint16)[pos].tolist()}


def main():
    _stub_bnb()
    torch.manual_seed(0)
    meta: dict = {"generator": "tests/golden/make_golden.py", "reference": "toddllm/gpt-oss-vllm @ /root/reference"}
    arrays: dict = {}

    sb = _load("streaming_unsloth_bnb")
    dl = _load("unsloth_direct_loader")
    cv = _load("unsloth_to_vllm_converter")

    # NF4 table as the reference holds it (P/streaming_unsloth_bnb.py:101-106)
    arrays["nf4_quant_map"] = sb.NF4_QUANT_MAP.numpy()

    # 1) dequant_unsloth_expert_cpu on full-size synthetic experts (shapes are hard-coded in the reference)
    with tempfile.TemporaryDirectory() as td:
        for name, proj, n, scalar, seed in (("gate_up_vec", "gate_up", 5760 * 2880, False, 101),
                                            ("down_scalar", "down", 2880 * 2880, True, 102)):
            st = synth_state(seed, n, scalar)
            p = os.path.join(td, name + ".safetensors")
            write_ckpt(p, st, 0, 3, proj + "_projs")
            out = sb.dequant_unsloth_expert_cpu([p], 0, 3, proj)
            assert out is not None and out.dtype == torch.float16
            meta["dequant_unsloth_expert_cpu/" + name] = {"seed": seed, "n_weights": n, "scalar_nested": scalar,
                                                          **digest(out.numpy())}
        # 2) unsloth_direct_loader.load_and_dequantize_unsloth_weight (per-block python loop) on 'down'
        st = synth_state(103, 2880 * 2880, False)
        p = os.path.join(td, "direct.safetensors")
        write_ckpt(p, st, 1, 5, "down_projs")
        out = dl.load_and_dequantize_unsloth_weight(p, 1, 5, "down")
        meta["load_and_dequantize_unsloth_weight/down_vec"] = {"seed": 103, "n_weights": 2880 * 2880,
                                                               "scalar_nested": False, **digest(out.numpy())}

    # 3) dequantize_absmax, both branches (small)
    rng = np.random.default_rng(7)
    u8 = rng.integers(0, 256, size=1024, dtype=np.uint8)
    qmap = orc.create_dynamic_map()
    for tag, nested in (("scalar", np.array([0.037], np.float32)),
                        ("vec", (rng.random(4, dtype=np.float32) + np.float32(0.5)))):
        out = dl.dequantize_absmax(torch.from_numpy(u8), torch.from_numpy(nested), torch.from_numpy(qmap))
        arrays[f"dequantize_absmax/{tag}/u8"] = u8
        arrays[f"dequantize_absmax/{tag}/nested"] = nested
        arrays[f"dequantize_absmax/{tag}/out"] = out.numpy()

    # 4) UnslothMoEWrapper.forward with plain fp32 nn.Linear experts (the wrapper only calls module(x))
    E, H, I, B, S = 8, 48, 40, 2, 5
    g = torch.Generator().manual_seed(11)
    w13 = [torch.nn.Linear(H, 2 * I, bias=False) for _ in range(E)]
    w2 = [torch.nn.Linear(I, H, bias=False) for _ in range(E)]
    for m in w13 + w2:
        m.weight.data = torch.randn(m.weight.shape, generator=g) * 0.2
    wrap = cv.UnslothMoEWrapper({"w13": w13, "w2": w2}, num_experts=E)
    xs = torch.randn(B, S, H, generator=g)
    lg = torch.randn(B, S, E, generator=g)
    out = wrap.forward(xs, lg, top_k=4)
    arrays["wrapper/x"] = xs.numpy()
    arrays["wrapper/logits"] = lg.numpy()
    arrays["wrapper/w13"] = torch.stack([m.weight.data for m in w13]).numpy()
    arrays["wrapper/w2"] = torch.stack([m.weight.data for m in w2]).numpy()
    arrays["wrapper/out"] = out.numpy()

    # 5) the reference's two routing lines verbatim (P/unsloth_to_vllm_converter.py:291-292), fp16 + ties
    lg16 = (torch.randn(64, 32, generator=g) * 2).half()
    lg16[3, 5] = lg16[3, 9] = lg16[3].max() + 1      # exact tie for first place
    lg16[7, :] = 0.5                                  # fully tied row
    tw, ti = torch.topk(lg16.float(), 4, dim=-1)
    tw = torch.nn.functional.softmax(tw, dim=-1)
    arrays["route/logits_f16"] = lg16.numpy()
    arrays["route/ids"] = ti.numpy().astype(np.int32)
    arrays["route/weights"] = tw.numpy()

    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **arrays)
    with open(os.path.join(HERE, "reference_golden.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("wrote", len(arrays), "arrays,", len(meta), "meta entries")


if __name__ == "__main__":
    main()
