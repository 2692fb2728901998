"""Bind a pre-quantised bitsandbytes-NF4 gpt-oss checkpoint (the Unsloth `*-bnb-4bit` safetensors the reference works from)
to `NF4ExpertWeights` WITHOUT unpacking or re-quantising ("NEVER unpack qweight", P/unsloth_bnb_analysis.md:57).  SURVEY 8(f).1.

On-disk key set per expert matrix (P/unsloth_bnb_loader.py:196-227, P/debug_shapes.py:31-46, dtypes P/unsloth_direct_loader.py:69-82):

    model.layers.{L}.mlp.experts.gate_up_projs.{e}.weight                    uint8 [rows*cols/2, 1]   packed NF4, high nibble first
    ....weight.absmax                                                        uint8 [rows*cols/64]     double-quantised absmax
    ....weight.nested_absmax                                                 fp32  [ceil(nb/256)]
    ....weight.nested_quant_map                                              fp32  [256]              (code2)
    ....weight.quant_map                                                     fp32  [16]               (must be the NF4 table)
    ....weight.quant_state.bitsandbytes__nf4                                 uint8 [n]                JSON: nested_offset, blocksize, ...
    (same under down_projs.{e})

The reference's own loaders dequantise these on the CPU in Python loops (P/unsloth_direct_loader.py:51-142, "15-20 minutes",
P/README.md:98) and then drop the 4-bit form (P/patch_11_consolidate_unsloth.py:48-69).  Here the bytes are stacked expert-major
into the fused layout of SURVEY 8 and handed to the kernels as they are; the nested absmax is kept nested.

This module is host-side plumbing (safetensors -> torch tensors); it does no arithmetic on the weights.
"""
from __future__ import annotations

import json
from typing import Callable, Iterable

import torch

from .quant import BLOCKSIZE, NESTED_BLOCKSIZE, NF4ExpertWeights

# P/streaming_unsloth_bnb.py:101-106 -- used only to VERIFY the checkpoint's quant_map; the kernels carry their own copy
NF4_TABLE = (-1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
             -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
             0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941, 0.7229568362236023, 1.0)

PROJ_KEYS = {"w13": "gate_up_projs", "w2": "down_projs"}


class CheckpointFormatError(ValueError):
    pass


def expert_prefix(layer: int, proj: str, expert: int) -> str:
    return f"model.layers.{layer}.mlp.experts.{PROJ_KEYS[proj]}.{expert}.weight"


def _offset_from_state_blob(blob: torch.Tensor) -> tuple[float, dict]:
    """bitsandbytes packs the non-tensor part of a QuantState as JSON bytes; `nested_offset` is the double-quant offset."""
    meta = json.loads(bytes(blob.to(torch.uint8).flatten().tolist()).decode("utf-8"))
    if "nested_offset" not in meta:
        raise CheckpointFormatError("quant_state blob has no nested_offset (not a double-quantised state)")
    return float(meta["nested_offset"]), meta


def load_expert_matrix(get: Callable[[str], torch.Tensor | None], prefix: str, rows: int, cols: int,
                       offset_override: float | None = None) -> dict:
    """One expert matrix as the quant-state dict `NF4ExpertWeights.from_states` takes.  `get(key)` returns a tensor or None."""
    n = rows * cols
    nb = n // BLOCKSIZE
    n2 = -(-nb // NESTED_BLOCKSIZE)
    packed = get(prefix)
    if packed is None:
        raise CheckpointFormatError(f"missing tensor {prefix}")      # the reference silently zero-filled (P/streaming_unsloth_bnb.py:261-264)
    if packed.dtype != torch.uint8 or packed.numel() != n // 2:
        raise CheckpointFormatError(f"{prefix}: expected {n // 2} packed uint8 bytes for [{rows}, {cols}], got "
                                    f"{tuple(packed.shape)} {packed.dtype}")
    absmax = get(prefix + ".absmax")
    absmax2 = get(prefix + ".nested_absmax")
    code2 = get(prefix + ".nested_quant_map")
    qmap = get(prefix + ".quant_map")
    if absmax is None or absmax2 is None or code2 is None:
        raise CheckpointFormatError(f"{prefix}: missing .absmax / .nested_absmax / .nested_quant_map")
    if absmax.dtype != torch.uint8 or absmax.numel() != nb:
        raise CheckpointFormatError(f"{prefix}.absmax: expected {nb} uint8 (blocksize 64), got {tuple(absmax.shape)} {absmax.dtype}")
    if absmax2.numel() != n2:
        raise CheckpointFormatError(f"{prefix}.nested_absmax: expected {n2} values (nested blocksize 256), got {absmax2.numel()}")
    if code2.numel() != 256:
        raise CheckpointFormatError(f"{prefix}.nested_quant_map: expected 256 values")
    if qmap is not None:
        if qmap.numel() != 16 or not torch.equal(qmap.flatten().to(torch.float32), torch.tensor(NF4_TABLE, dtype=torch.float32)):
            raise CheckpointFormatError(f"{prefix}.quant_map is not the NF4 table: this path is NF4 only")
    if offset_override is not None:
        offset = float(offset_override)
    else:
        blob = get(prefix + ".quant_state.bitsandbytes__nf4")
        if blob is None:
            raise CheckpointFormatError(f"{prefix}: no quant_state blob to read nested_offset from; pass offset_override=")
        offset, meta = _offset_from_state_blob(blob)
        if int(meta.get("blocksize", BLOCKSIZE)) != BLOCKSIZE or int(meta.get("nested_blocksize", NESTED_BLOCKSIZE)) != NESTED_BLOCKSIZE:
            raise CheckpointFormatError(f"{prefix}: blocksize {meta.get('blocksize')} / nested {meta.get('nested_blocksize')} unsupported")
        if meta.get("quant_type", "nf4") != "nf4":
            raise CheckpointFormatError(f"{prefix}: quant_type {meta.get('quant_type')} unsupported")
    return {"packed": packed.reshape(-1).numpy(), "absmax_u8": absmax.reshape(-1).numpy(),
            "absmax2": absmax2.reshape(-1).to(torch.float32).numpy(), "code2": code2.reshape(-1).to(torch.float32).numpy(),
            "offset": offset, "shape": (rows, cols), "blocksize": BLOCKSIZE}


def load_layer_experts(files: Iterable[str], layer: int, num_experts: int, hidden_size: int, intermediate_size: int,
                       device="cpu", experts: range | None = None) -> tuple[NF4ExpertWeights, NF4ExpertWeights]:
    """(w13, w2) of one MoE layer, fused expert-major (`[E, 2I, H]` / `[E, H, I]` packed), nested absmax kept.
    `experts` selects a sub-range (expert parallelism: each rank loads only its own experts)."""
    from safetensors import safe_open
    handles = [safe_open(f, framework="pt", device="cpu") for f in files]
    index = {}
    for h in handles:
        for k in h.keys():
            index[k] = h

    def get(key):
        h = index.get(key)
        return None if h is None else h.get_tensor(key)

    ids = experts if experts is not None else range(num_experts)
    st13 = [load_expert_matrix(get, expert_prefix(layer, "w13", e), 2 * intermediate_size, hidden_size) for e in ids]
    st2 = [load_expert_matrix(get, expert_prefix(layer, "w2", e), hidden_size, intermediate_size) for e in ids]
    for sts in (st13, st2):   # one shared code2 table per matrix set is the layout the kernels take (code2_stride = 0)
        for s in sts[1:]:
            if not (s["code2"] == sts[0]["code2"]).all():
                raise CheckpointFormatError("experts of one layer carry different nested_quant_map tables")
    return NF4ExpertWeights.from_states(st13, device), NF4ExpertWeights.from_states(st2, device)


def save_layer_experts(path: str, layer: int, st13: list[dict], st2: list[dict]) -> None:
    """Write quant-state dicts in the same on-disk key set (used by tests and to make synthetic checkpoints)."""
    from safetensors.torch import save_file
    out = {}
    for proj, sts in (("w13", st13), ("w2", st2)):
        for e, s in enumerate(sts):
            p = expert_prefix(layer, proj, e)
            rows, cols = s["shape"]
            out[p] = torch.from_numpy(s["packed"].copy()).reshape(-1, 1)
            out[p + ".absmax"] = torch.from_numpy(s["absmax_u8"].copy())
            out[p + ".nested_absmax"] = torch.from_numpy(s["absmax2"].copy())
            out[p + ".nested_quant_map"] = torch.from_numpy(s["code2"].copy())
            out[p + ".quant_map"] = torch.tensor(NF4_TABLE, dtype=torch.float32)
            blob = json.dumps({"quant_type": "nf4", "blocksize": BLOCKSIZE, "nested_blocksize": NESTED_BLOCKSIZE,
                               "nested_offset": float(s["offset"]), "shape": [rows, cols], "dtype": "float16"}).encode()
            out[p + ".quant_state.bitsandbytes__nf4"] = torch.tensor(list(blob), dtype=torch.uint8)
    save_file(out, path)
