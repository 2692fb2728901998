"""CPU oracle for the bitsandbytes-NF4 fused-MoE expert path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this module; the product path (``gpt-oss-vllm_b200/``) never does and fails loudly without its CUDA library.

What it restates (R/ = /root/reference/, P/ = R/patches/, V/ = the vLLM 0.22 in this image):

* NF4 code table ............... P/streaming_unsloth_bnb.py:101-106 (16 fp32 constants, verbatim values)
* blockwise scale structure .... P/streaming_unsloth_bnb.py:199-202 (``view(-1, blocksize) * absmax.unsqueeze(1)``)
* double-quant absmax formula .. V/model_executor/model_loader/bitsandbytes_loader.py:599-601
                                 (``dequantize_blockwise(absmax_u8, state2) + offset``)
* routing ...................... P/unsloth_to_vllm_converter.py:291-292 (top-k on logits, softmax over the k)
* per-(token,k) expert loop .... P/unsloth_to_vllm_converter.py:298-323
* gpt-oss activation ........... V/model_executor/layers/activation.py:389-397 (interleaved clamped SwiGLU)
* reference activation ......... P/unsloth_to_vllm_converter.py:312-315 (``chunk`` + ``silu``) -- divergence mode

PARITY PINNING STATUS:
  * pinned against the reference's OWN Python, run in the build container with ``bitsandbytes`` stubbed
    (tests/golden/make_golden.py): the reference-variant dequant (`dequant_reference_variant`,
    `dequantize_absmax_reference`), the router + expert loop + chunk/silu forward (`moe_forward_ref(act="chunk_silu")`).
  * the *bnb-semantics* dequant (`dequant_nf4`: high nibble first, ``+ offset``) follows the bitsandbytes library,
    which is an unpinned, un-vendored dependency of the reference and absent from this image: **parity unpinned**
    for that one function -- it is anchored on the formula vLLM carries (V/...bitsandbytes_loader.py:599-601), the
    NF4 table, and the known-answer vectors in tests/test_oracle.py.
"""
from __future__ import annotations

import numpy as np

# P/streaming_unsloth_bnb.py:101-106 -- all 16 are exactly representable in fp32.
NF4_CODE = np.array(
    [
        -1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453,
        -0.28444138169288635, -0.18477343022823334, -0.09105003625154495, 0.0,
        0.07958029955625534, 0.16093020141124725, 0.24611230194568634, 0.33791524171829224,
        0.44070982933044434, 0.5626170039176941, 0.7229568362236023, 1.0,
    ],
    dtype=np.float32,
)

BLOCKSIZE = 64          # north-star; P/unsloth_bnb_loader.py:71,112
NESTED_BLOCKSIZE = 256  # bnb double-quant block (SURVEY §8 header)
SWIGLU_ALPHA = 1.702
SWIGLU_LIMIT = 7.0


# --------------------------------------------------------------------------------------------------------------
# 8-bit dynamic map used by bnb for the nested absmax (data, not algorithm: it travels in the quant state).
# --------------------------------------------------------------------------------------------------------------
def create_dynamic_map(signed: bool = True, max_exponent_bits: int = 7, total_bits: int = 8) -> np.ndarray:
    """256 sorted fp32 values: bnb's ``create_dynamic_map`` recipe (recalled; SURVEY §8c says any sorted table
    works because oracle and kernel read the same array)."""
    data: list[float] = []
    non_sign_bits = total_bits - 1
    additional_items = 2 ** (non_sign_bits - max_exponent_bits) - 1
    i = 0
    for i in range(max_exponent_bits):
        fraction_items = int(
            2 ** (i + non_sign_bits - max_exponent_bits) + 1
            if signed
            else 2 ** (i + non_sign_bits - max_exponent_bits + 1) + 1
        )
        boundaries = np.linspace(0.1, 1, fraction_items, dtype=np.float32)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
    if additional_items > 0:
        boundaries = np.linspace(0.1, 1, additional_items + 1, dtype=np.float32)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
    data.append(0.0)
    data.append(1.0)
    data += [0.0] * (2 ** total_bits - len(data))
    data.sort()
    return np.asarray(data, dtype=np.float32)


# --------------------------------------------------------------------------------------------------------------
# Quantiser (needed only to make weights locally; dequant is what is graded bit-exact).
# Follows bnb ``quantize_4bit(blocksize=64, compress_statistics=True, quant_type='nf4')`` as invoked at
# V/model_executor/model_loader/bitsandbytes_loader.py:433-438 and P/bnb_moe_4bit.py:86-91.
# --------------------------------------------------------------------------------------------------------------
def _nearest_code(values: np.ndarray, code_sorted: np.ndarray) -> np.ndarray:
    """Index of the nearest entry of a sorted code table (ties -> lower index)."""
    mid = (code_sorted[:-1].astype(np.float64) + code_sorted[1:].astype(np.float64)) / 2.0
    return np.searchsorted(mid, values.astype(np.float64), side="left").astype(np.uint8)


def quantize_nf4(w: np.ndarray, blocksize: int = BLOCKSIZE, nested_blocksize: int = NESTED_BLOCKSIZE,
                 code2: np.ndarray | None = None) -> dict:
    """Quantise a [out, in] matrix (row-major flat) to packed NF4 with double-quantised absmax.

    Returns a dict mirroring bnb's ``QuantState`` fields and the Unsloth on-disk key set
    (P/unsloth_bnb_loader.py:202-222): ``packed`` u8[N/2] (high nibble = even element), ``absmax_u8`` u8[nb],
    ``absmax2`` f32[ceil(nb/256)] (= ``nested_absmax``), ``code2`` f32[256] (= ``nested_quant_map``),
    ``offset`` f32 scalar, ``code`` f32[16] (= ``quant_map``), ``shape``, ``blocksize``.
    ``absmax_f32`` is the DE-NESTED absmax the runtime actually uses (what vLLM keeps after
    ``_dequantize_dq``, V/...bitsandbytes_loader.py:585-617).
    """
    shape = tuple(w.shape)
    flat = np.ascontiguousarray(w, dtype=np.float32).reshape(-1)
    n = flat.size
    assert n % blocksize == 0 and n % 2 == 0, "path shapes are multiples of 64"
    blocks = flat.reshape(-1, blocksize)
    absmax = np.abs(blocks).max(axis=1).astype(np.float32)
    safe = np.where(absmax > 0, absmax, np.float32(1.0))
    normed = blocks / safe[:, None]
    q = _nearest_code(normed.reshape(-1), NF4_CODE)
    packed = ((q[0::2] << 4) | q[1::2]).astype(np.uint8)

    if code2 is None:
        code2 = create_dynamic_map()
    offset = np.float32(absmax.mean(dtype=np.float64))
    centred = (absmax - offset).astype(np.float32)
    nb = centred.size
    n2 = -(-nb // nested_blocksize)
    padded = np.zeros(n2 * nested_blocksize, dtype=np.float32)
    padded[:nb] = centred
    pb = padded.reshape(n2, nested_blocksize)
    absmax2 = np.abs(pb).max(axis=1).astype(np.float32)
    safe2 = np.where(absmax2 > 0, absmax2, np.float32(1.0))
    absmax_u8 = _nearest_code((pb / safe2[:, None]).reshape(-1), code2)[:nb]
    st = {
        "packed": packed, "absmax_u8": absmax_u8, "absmax2": absmax2, "code2": code2.astype(np.float32),
        "offset": offset, "code": NF4_CODE.copy(), "shape": shape, "blocksize": blocksize,
        "nested_blocksize": nested_blocksize,
    }
    st["absmax_f32"] = dequant_absmax_nested(absmax_u8, absmax2, st["code2"], offset, nested_blocksize)
    return st


# --------------------------------------------------------------------------------------------------------------
# Dequant -- bnb semantics (the parity target).
# --------------------------------------------------------------------------------------------------------------
def dequant_absmax_nested(absmax_u8: np.ndarray, absmax2: np.ndarray, code2: np.ndarray, offset,
                          nested_blocksize: int = NESTED_BLOCKSIZE) -> np.ndarray:
    """``absmax[b] = fp32(code2[absmax_u8[b]] * absmax2[b // 256]) + offset`` -- multiply and add separately
    rounded to fp32 (no FMA).  V/model_executor/model_loader/bitsandbytes_loader.py:599-601."""
    idx = np.arange(absmax_u8.size) // nested_blocksize
    prod = (code2.astype(np.float32)[absmax_u8.astype(np.int64)] * absmax2.astype(np.float32)[idx]).astype(np.float32)
    return (prod + np.float32(offset)).astype(np.float32)


def unpack_nibbles(packed: np.ndarray, nibble_order: str = "high_first") -> np.ndarray:
    """bnb: element 2j = byte>>4, element 2j+1 = byte&0xF.  ``low_first`` is the reference's own (divergent)
    order, P/streaming_unsloth_bnb.py:181-187."""
    packed = packed.reshape(-1)
    q = np.empty(packed.size * 2, dtype=np.uint8)
    if nibble_order == "high_first":
        q[0::2] = packed >> 4
        q[1::2] = packed & 0x0F
    elif nibble_order == "low_first":
        q[0::2] = packed & 0x0F
        q[1::2] = packed >> 4
    else:
        raise ValueError(nibble_order)
    return q


def dequant_nf4(packed: np.ndarray, absmax_f32: np.ndarray, shape, blocksize: int = BLOCKSIZE,
                nibble_order: str = "high_first", code: np.ndarray = NF4_CODE) -> np.ndarray:
    """``w16[i] = fp16_rn( fp32(code[q_i]) * fp32(absmax[i // 64]) )``: one fp32 multiply, one RN conversion.
    Structure: P/streaming_unsloth_bnb.py:199-202, 228-229; nibble order and fp32-then-fp16 rounding: bnb."""
    q = unpack_nibbles(packed, nibble_order)
    vals = code.astype(np.float32)[q].reshape(-1, blocksize)
    out = (vals * absmax_f32.astype(np.float32)[:, None]).astype(np.float32)
    return out.astype(np.float16).reshape(shape)


def dequant_state(st: dict, nested: bool = True) -> np.ndarray:
    absmax = (dequant_absmax_nested(st["absmax_u8"], st["absmax2"], st["code2"], st["offset"],
                                    st.get("nested_blocksize", NESTED_BLOCKSIZE))
              if nested else st["absmax_f32"])
    return dequant_nf4(st["packed"], absmax, st["shape"], st["blocksize"])


# --------------------------------------------------------------------------------------------------------------
# Dequant -- the reference's own Python variants (named divergences; pinned by tests/golden).
# --------------------------------------------------------------------------------------------------------------
def dequantize_absmax_reference(absmax_u8: np.ndarray, nested_absmax: np.ndarray, nested_quant_map: np.ndarray
                                ) -> np.ndarray:
    """P/unsloth_direct_loader.py:21-48: ``map[u8] * nested_absmax`` with NO ``+offset``; for a vector
    ``nested_absmax`` the block size is ``len(indices) // len(nested_absmax)`` (floor -- 255 for the real
    259,200/1,013 geometry, i.e. *not* bnb's 256) and elements beyond ``len(nested)*blocksize`` stay unscaled."""
    deq = nested_quant_map.astype(np.float32)[absmax_u8.reshape(-1).astype(np.int64)].copy()
    if nested_absmax.size == 1:
        # `.item()` promotes to a python double; torch multiplies the fp32 tensor by it in fp32.
        return (deq * np.float32(nested_absmax.reshape(-1)[0])).astype(np.float32)
    bs = deq.size // nested_absmax.size
    for i in range(nested_absmax.size):
        deq[i * bs:min((i + 1) * bs, deq.size)] *= np.float32(nested_absmax[i])
    return deq


def dequant_reference_variant(packed: np.ndarray, absmax_u8: np.ndarray, nested_absmax: np.ndarray,
                              nested_quant_map: np.ndarray, shape) -> np.ndarray:
    """P/streaming_unsloth_bnb.py:148-229 as it behaves on the real Unsloth geometry: low-nibble-first unpack
    (:181-187), absmax = ``map[u8] * nested_absmax`` without offset where a shape mismatch falls back to
    ``nested_absmax.mean()`` (:158-171), blocksize inferred as ``numel // absmax.numel()`` (:196-202)."""
    absmax = nested_quant_map.astype(np.float32)[absmax_u8.reshape(-1).astype(np.int64)]
    if nested_absmax.size == 1:
        absmax = (absmax * np.float32(nested_absmax.reshape(-1)[0])).astype(np.float32)
    elif nested_absmax.shape == absmax.shape:
        absmax = (absmax * nested_absmax.astype(np.float32)).astype(np.float32)
    else:
        import torch  # mean() must match torch's fp32 reduction order bit-for-bit
        m = torch.from_numpy(np.ascontiguousarray(nested_absmax, dtype=np.float32)).mean()
        absmax = (torch.from_numpy(absmax.copy()) * m).numpy()
    n = int(np.prod(shape))
    q = unpack_nibbles(packed, "low_first")[:n]
    vals = NF4_CODE[q]
    assert n % absmax.size == 0
    bs = n // absmax.size
    out = (vals.reshape(-1, bs) * absmax[:, None]).astype(np.float32)
    return out.astype(np.float16).reshape(shape)


# --------------------------------------------------------------------------------------------------------------
# Router, activation, MoE forward.
# --------------------------------------------------------------------------------------------------------------
def route_topk(logits: np.ndarray, k: int) -> tuple[np.ndarray, np.ndarray]:
    """P/unsloth_to_vllm_converter.py:291-292 == softmax -> top-k -> renormalise
    (V/model_executor/layers/fused_moe/router/fused_topk_router.py:69-102, renormalize=True).
    fp32 math; ids in descending-logit order, ties -> lowest index.  Returns (ids int32 [T,k], w fp32 [T,k])."""
    l32 = logits.astype(np.float32)
    ids = np.argsort(-l32, axis=-1, kind="stable")[:, :k].astype(np.int32)
    sel = np.take_along_axis(l32, ids.astype(np.int64), axis=-1)
    e = np.exp(sel - sel.max(axis=-1, keepdims=True), dtype=np.float32)
    w = (e / e.sum(axis=-1, keepdims=True, dtype=np.float32)).astype(np.float32)
    return ids, w


def swiglu_oai(gu: np.ndarray, alpha: float = SWIGLU_ALPHA, limit: float = SWIGLU_LIMIT) -> np.ndarray:
    """V/model_executor/layers/activation.py:389-397: gate = x[..., ::2] clamped to <= limit, up = x[..., 1::2]
    clamped to [-limit, limit], out = (up + 1) * gate * sigmoid(alpha * gate).  fp32 in, fp32 out."""
    g = np.minimum(gu[..., 0::2].astype(np.float32), np.float32(limit))
    u = np.clip(gu[..., 1::2].astype(np.float32), np.float32(-limit), np.float32(limit))
    glu = g * (np.float32(1.0) / (np.float32(1.0) + np.exp(-np.float32(alpha) * g, dtype=np.float32)))
    return ((u + np.float32(1.0)) * glu).astype(np.float32)


def chunk_silu(gu: np.ndarray) -> np.ndarray:
    """P/unsloth_to_vllm_converter.py:312-315: ``gate, up = chunk(2)``; ``silu(gate) * up``."""
    half = gu.shape[-1] // 2
    g = gu[..., :half].astype(np.float32)
    u = gu[..., half:].astype(np.float32)
    return (g / (np.float32(1.0) + np.exp(-g, dtype=np.float32)) * u).astype(np.float32)


def _mm_f16_f32acc(a16: np.ndarray, b16_t: np.ndarray) -> np.ndarray:
    """fp16 operands, fp32 accumulate (products of two fp16 are exact in fp32), fp32 result."""
    import torch
    return (torch.from_numpy(a16.astype(np.float32)) @ torch.from_numpy(b16_t.astype(np.float32)).T).numpy()


def moe_forward_ref(x: np.ndarray, logits: np.ndarray, w13: np.ndarray, w2: np.ndarray, k: int = 4,
                    act: str = "swiglu_oai", topk: tuple[np.ndarray, np.ndarray] | None = None,
                    storage=np.float16, b13: np.ndarray | None = None, b2: np.ndarray | None = None) -> np.ndarray:
    """The MoE layer of P/unsloth_to_vllm_converter.py:282-325, grouped by expert instead of by token
    (same arithmetic per (token, k) pair): ``h = x @ W13_e^T`` (fp32 acc -> storage) -> activation in fp32 ->
    storage -> ``y = h @ W2_e^T`` (fp32 acc -> storage) -> ``out[t] += w[t,k] * y`` in fp32 (k ascending)
    -> storage.  ``x`` [T,H], ``logits`` [T,E], ``w13`` [E,2I,H], ``w2`` [E,H,I] dequantised.
    ``storage=np.float32`` gives the all-fp32 variant used to pin against the reference run with fp32 experts."""
    T, H = x.shape
    ids, w = route_topk(logits, k) if topk is None else topk
    y_pair = np.zeros((T, k, H), dtype=np.float32)
    fn = swiglu_oai if act == "swiglu_oai" else chunk_silu
    for e in np.unique(ids):
        tt, kk = np.nonzero(ids == e)
        xe = x[tt].astype(storage)
        # expert biases (gpt-oss, V/model_executor/models/gpt_oss.py:199; not in the reference wrapper): added to the fp32
        # accumulator of each linear layer before its one rounding to storage, as F.linear(x, W, b) does
        h = _mm_f16_f32acc(xe, w13[e])
        if b13 is not None:
            h = h + b13[e].astype(np.float32)
        h = h.astype(storage)
        a = fn(h.astype(np.float32)).astype(storage)
        y = _mm_f16_f32acc(a, w2[e])
        if b2 is not None:
            y = y + b2[e].astype(np.float32)
        y = y.astype(storage)
        y_pair[tt, kk] = y.astype(np.float32)
    out = np.zeros((T, H), dtype=np.float32)
    for j in range(k):
        out += w[:, j:j + 1].astype(np.float32) * y_pair[:, j]
    return out.astype(storage)


def moe_forward_from_states(x, logits, st13: list[dict], st2: list[dict], k: int = 4, nested: bool = True,
                            act: str = "swiglu_oai", topk=None, b13=None, b2=None) -> np.ndarray:
    """The reference path end to end: bnb-semantics dequant of every expert, then `moe_forward_ref`."""
    w13 = np.stack([dequant_state(s, nested) for s in st13])
    w2 = np.stack([dequant_state(s, nested) for s in st2])
    return moe_forward_ref(x, logits, w13, w2, k, act, topk, b13=b13, b2=b2)
