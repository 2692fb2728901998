"""Packed NF4 expert weights in the reference's layout, and the device (de)quantiser entry points.

Layout (unchanged from bitsandbytes / vLLM, SURVEY 8 header):
  packed     uint8 [E, rows*cols/2, 1]   element 2j in the high nibble       (V/.../quantization/bitsandbytes.py:508-575)
  flat form  absmax fp32 [E, rows*cols/64]                                  (vLLM after _dequantize_dq)
  nested     absmax_u8 [E, nb], absmax2 fp32 [E, ceil(nb/256)], code2 fp32 [256], offset fp32 [E]
             == the on-disk key set {.absmax, .nested_absmax, .nested_quant_map}  (P/unsloth_bnb_loader.py:202-222)
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib

BLOCKSIZE = 64
NESTED_BLOCKSIZE = 256


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: this package has no CPU fallback")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


_CODE2_TABLES: dict = {}


def _shared_code2(t: torch.Tensor | None) -> torch.Tensor | None:
    """One device copy per distinct 256-entry double-quant table: w13 and w2 of a layer (of every layer, in practice: bnb's
    dynamic map) then hand the SAME pointer to the library, which is what lets the merged w13+w2 decode launch park the table in
    its decode table's holes (csrc/gemm_dec.cuh, dec_launch_sides)."""
    if t is None or not t.is_cuda or t.dim() != 1 or t.numel() != 256 or t.dtype != torch.float32:
        return t
    key = (t.device.index, t.detach().cpu().numpy().tobytes())
    return _CODE2_TABLES.setdefault(key, t)


def dynamic_code_map() -> torch.Tensor:
    """The 256-entry table (`code2`, fp32, ascending) that double quantisation stores the absmax bytes against: bitsandbytes'
    signed 8-bit "dynamic" map as used by quantize_4bit(compress_statistics=True) (V/model_executor/model_loader/
    bitsandbytes_loader.py:433-438 is where vLLM triggers it).  Seven decades 1e-6 .. 1e0; decade d carries 2^d magnitudes --
    the midpoints of 2^d equal steps across [0.1, 1] -- with both signs, plus exactly 0 and exactly 1: 2 * 127 + 2 = 256 values.
    It is DATA as far as dequantisation goes (the table travels in the quant state), so this exists only for callers that
    quantise locally (NF4ExpertWeights.quantize, bench.py, tests)."""
    import numpy as np
    vals = [np.float32(0.0), np.float32(1.0)]
    for d in range(7):
        edges = np.linspace(0.1, 1, 2 ** d + 1, dtype=np.float32)
        mags = (10 ** (d - 6)) * ((edges[:-1] + edges[1:]) / 2.0)
        vals.extend(mags)
        vals.extend(-mags)
    table = np.sort(np.asarray(vals, dtype=np.float32))
    assert table.shape == (256,)
    return torch.from_numpy(table)



@dataclass
class NF4ExpertWeights:
    """One fused expert matrix set: E experts x [rows, cols]."""
    packed: torch.Tensor
    rows: int
    cols: int
    absmax: torch.Tensor | None = None      # flat fp32 [E, nb]
    absmax_u8: torch.Tensor | None = None   # nested
    absmax2: torch.Tensor | None = None
    code2: torch.Tensor | None = None
    offset: torch.Tensor | None = None

    def __post_init__(self):
        if self.cols % BLOCKSIZE:
            raise ValueError("cols must be a multiple of 64")
        if self.packed.dtype != torch.uint8:
            raise TypeError("packed must be uint8")
        self.num_experts = int(self.packed.shape[0])
        if self.packed.numel() != self.num_experts * self.rows * self.cols // 2:
            raise ValueError("packed has the wrong number of bytes for [E, rows, cols]")
        if self.absmax is None and not self.has_nested:
            raise ValueError("give a flat absmax or the nested state")
        self.code2 = _shared_code2(self.code2)

    @property
    def has_nested(self) -> bool:
        return all(t is not None for t in (self.absmax_u8, self.absmax2, self.code2, self.offset))

    @property
    def blocks_per_expert(self) -> int:
        return self.rows * self.cols // BLOCKSIZE

    @property
    def device(self):
        return self.packed.device

    def nbytes(self, nested: bool) -> int:
        """Algorithmic bytes per expert (packed + the absmax form the kernels read)."""
        nb = self.blocks_per_expert
        a = nb + 4 * (-(-nb // NESTED_BLOCKSIZE)) if nested else 4 * nb
        return self.rows * self.cols // 2 + a

    def c_struct(self, nested: bool | None = None) -> _lib.QWeight:
        """struct nf4moe_qweight for the C ABI.  nested=None picks flat when present."""
        if nested is None:
            nested = self.absmax is None
        if nested and not self.has_nested:
            raise ValueError("nested state not available")
        if not nested and self.absmax is None:
            raise ValueError("flat absmax not available (call denest())")
        q = _lib.QWeight()
        q.packed = self.packed.data_ptr()
        q.absmax_f32 = None if nested else self.absmax.data_ptr()
        q.absmax_u8 = _ptr(self.absmax_u8) if nested else None
        q.absmax2 = _ptr(self.absmax2) if nested else None
        q.code2 = _ptr(self.code2) if nested else None
        q.offset = _ptr(self.offset) if nested else None
        q.code2_stride = 0 if (self.code2 is None or self.code2.dim() == 1) else 256
        q.rows, q.cols = self.rows, self.cols
        return q

    # ---- constructors -------------------------------------------------------------------------------------
    @classmethod
    def quantize(cls, w: torch.Tensor, code2: torch.Tensor) -> "NF4ExpertWeights":
        """fp16 [E, rows, cols] on the GPU -> packed NF4 with double-quantised absmax (nf4moe_quantize)."""
        _require_cuda(w, "w")
        if w.dtype != torch.float16 or w.dim() != 3:
            raise TypeError("w must be fp16 [E, rows, cols]")
        lib = _lib.load_library()
        E, rows, cols = w.shape
        nb = rows * cols // BLOCKSIZE
        n2 = -(-nb // NESTED_BLOCKSIZE)
        dev = w.device
        packed = torch.empty(E, rows * cols // 2, 1, dtype=torch.uint8, device=dev)
        absmax_u8 = torch.empty(E, nb, dtype=torch.uint8, device=dev)
        absmax2 = torch.empty(E, n2, dtype=torch.float32, device=dev)
        offset = torch.empty(E, dtype=torch.float32, device=dev)
        absmax = torch.empty(E, nb, dtype=torch.float32, device=dev)
        code2 = code2.to(device=dev, dtype=torch.float32).contiguous()
        scratch = torch.empty(E * nb * 4 + E * 8 + 512, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(lib.nf4moe_quantize(w.data_ptr(), E, rows, cols, packed.data_ptr(), absmax_u8.data_ptr(),
                                           absmax2.data_ptr(), code2.data_ptr(), offset.data_ptr(), absmax.data_ptr(),
                                           scratch.data_ptr(), _stream_ptr(dev)))
        return cls(packed, rows, cols, absmax, absmax_u8, absmax2, code2, offset)

    @classmethod
    def from_states(cls, states: list[dict], device) -> "NF4ExpertWeights":
        """From per-expert quant-state dicts (the oracle's / the Unsloth on-disk key set), without re-quantising."""
        import numpy as np
        rows, cols = states[0]["shape"]
        t = lambda k, dt: torch.from_numpy(np.stack([np.asarray(s[k]) for s in states])).to(device=device, dtype=dt)
        packed = t("packed", torch.uint8).reshape(len(states), -1, 1).contiguous()
        absmax = t("absmax_f32", torch.float32).contiguous() if "absmax_f32" in states[0] else None
        nested = {}
        if "absmax_u8" in states[0]:
            nested = dict(absmax_u8=t("absmax_u8", torch.uint8).contiguous(),
                          absmax2=t("absmax2", torch.float32).contiguous(),
                          code2=torch.from_numpy(np.asarray(states[0]["code2"], dtype=np.float32)).to(device),
                          offset=torch.tensor([float(s["offset"]) for s in states], dtype=torch.float32, device=device))
        return cls(packed, rows, cols, absmax, **nested)

    @classmethod
    def from_vllm_param(cls, param: torch.Tensor) -> "NF4ExpertWeights":
        """Bind vLLM's `w13_weight` / `w2_weight` (uint8 [E, n/2, 1]) and its fused `bnb_quant_state`
        (absmax already de-nested and fused expert-major, V/model_executor/model_loader/bitsandbytes_loader.py:619-696)."""
        qs = param.bnb_quant_state
        E, rows, cols = param.experts_shape
        if getattr(qs, "blocksize", BLOCKSIZE) != BLOCKSIZE or getattr(qs, "quant_type", "nf4") != "nf4":
            raise ValueError("only NF4 with blocksize 64 is supported on this path")
        absmax = qs.absmax
        if getattr(qs, "nested", False):
            raise ValueError("fused vLLM quant states are de-nested at load; got a nested one")
        return cls(param.data, rows, cols, absmax.to(torch.float32).reshape(E, -1).contiguous())

    # ---- operations ------------------------------------------------------------------------------------------
    def denest(self) -> torch.Tensor:
        """Fill (and return) the flat fp32 absmax from the nested state (nf4moe_dequant_absmax)."""
        lib = _lib.load_library()
        out = torch.empty(self.num_experts, self.blocks_per_expert, dtype=torch.float32, device=self.device)
        q = self.c_struct(nested=True)
        with torch.cuda.device(self.device):
            _lib.check(lib.nf4moe_dequant_absmax(C.byref(q), self.num_experts, out.data_ptr(), _stream_ptr(self.device)))
        self.absmax = out
        return out

    def dequantize(self, nested: bool | None = None, e0: int = 0, ne: int | None = None) -> torch.Tensor:
        """fp16 [ne, rows, cols], bit-exact bnb dequantize_4bit semantics (nf4moe_dequant)."""
        _require_cuda(self.packed, "packed")
        lib = _lib.load_library()
        ne = self.num_experts - e0 if ne is None else ne
        out = torch.empty(ne, self.rows, self.cols, dtype=torch.float16, device=self.device)
        q = self.c_struct(nested)
        with torch.cuda.device(self.device):
            _lib.check(lib.nf4moe_dequant(C.byref(q), e0, ne, out.data_ptr(), _stream_ptr(self.device)))
        return out
