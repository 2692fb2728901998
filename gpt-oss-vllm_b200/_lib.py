"""ctypes binding of libnf4moe.so (the C ABI declared in include/nf4moe.h).

There is no CPU or PyTorch fallback: if the library cannot be loaded the import of any compute entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnf4moe.so")

ABI_VERSION = 101          # NF4MOE_VERSION of include/nf4moe.h this binding was written against
ACT_SWIGLU_OAI = 0
ACT_CHUNK_SILU = 1
REGIME_AUTO, REGIME_GEMV, REGIME_GEMM, REGIME_GEMM_FAST = 0, 1, 2, 3
GEMM_EXACT, GEMM_FAST = 0, 1
LOGITS_DTYPE = {"float16": 0, "float32": 1, "bfloat16": 2}

# every symbol include/nf4moe.h declares (tests check the library exports exactly these)
EXPORTED = (
    "nf4moe_version", "nf4moe_auto_regime", "nf4moe_auto_regime_for", "nf4moe_cast_f16_bf16", "nf4moe_last_error", "nf4moe_workspace_bytes", "nf4moe_dequant", "nf4moe_dequant_absmax",
    "nf4moe_quantize", "nf4moe_quantize_static", "nf4moe_route_topk", "nf4moe_bucket", "nf4moe_permute_rows", "nf4moe_combine",
    "nf4moe_gemv_w13_swiglu", "nf4moe_gemv_w2", "nf4moe_gemm_w13_swiglu", "nf4moe_gemm_w2", "nf4moe_gemm_w13_w2", "nf4moe_moe_forward", "nf4moe_moe_forward_bias",
    "nf4moe_last_launch_count", "nf4moe_set_merged_decode", "nf4moe_ep_pack", "nf4moe_expert_rows", "nf4moe_combine_indexed",
    "nf4moe_workspace_status_offset", "nf4moe_expert_rows_bias", "nf4moe_combine_indexed_bias",
    "nf4moe_ep_dispatch_p2p", "nf4moe_ep_return_p2p", "nf4moe_ep_wait",
)
STATUS_EP_OVERFLOW, STATUS_DEC_NOFIT, STATUS_GEMM_NOFIT, STATUS_EP_TIMEOUT, STATUS_DEP_TIMEOUT = 1, 2, 4, 8, 16


class Nf4MoeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libnf4moe error {code}: {msg}")
        self.code = code


class QWeight(C.Structure):
    """struct nf4moe_qweight"""
    _fields_ = [
        ("packed", C.c_void_p), ("absmax_f32", C.c_void_p), ("absmax_u8", C.c_void_p), ("absmax2", C.c_void_p),
        ("code2", C.c_void_p), ("offset", C.c_void_p), ("code2_stride", C.c_int64), ("rows", C.c_int32),
        ("cols", C.c_int32),
    ]


_lock = threading.Lock()
_lib = None


def _declare(lib):
    vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
    qp = C.POINTER(QWeight)
    lib.nf4moe_version.restype = i32
    lib.nf4moe_auto_regime.restype = i32
    lib.nf4moe_auto_regime.argtypes = [i32] * 3
    lib.nf4moe_auto_regime_for.restype = i32
    lib.nf4moe_cast_f16_bf16.argtypes = [vp, vp, C.c_longlong, i32, vp]
    lib.nf4moe_auto_regime_for.argtypes = [i32] * 4
    lib.nf4moe_last_error.restype = C.c_char_p
    lib.nf4moe_last_launch_count.restype = i32
    lib.nf4moe_set_merged_decode.argtypes = [i32]
    lib.nf4moe_workspace_bytes.restype = sz
    lib.nf4moe_workspace_bytes.argtypes = [i32] * 5
    lib.nf4moe_workspace_status_offset.restype = sz
    lib.nf4moe_workspace_status_offset.argtypes = [i32] * 5
    lib.nf4moe_dequant.argtypes = [qp, i32, i32, vp, vp]
    lib.nf4moe_dequant_absmax.argtypes = [qp, i32, vp, vp]
    lib.nf4moe_quantize.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.nf4moe_quantize_static.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.nf4moe_route_topk.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp]
    lib.nf4moe_bucket.argtypes = [vp, i32, i32, i32, vp, vp, vp, vp, vp, i32, vp]
    lib.nf4moe_permute_rows.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
    lib.nf4moe_combine.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp, vp]
    lib.nf4moe_gemv_w13_swiglu.argtypes = [vp, qp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]
    lib.nf4moe_gemv_w2.argtypes = [vp, qp, vp, vp, vp, vp, i32, i32, i32, vp, vp]
    lib.nf4moe_gemm_w13_swiglu.argtypes = [vp, qp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]
    lib.nf4moe_gemm_w2.argtypes = [vp, qp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]
    lib.nf4moe_gemm_w13_w2.argtypes = [vp, qp, vp, vp, qp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    lib.nf4moe_moe_forward.argtypes = [vp, vp, i32, vp, vp, qp, qp, vp, i32, i32, i32, i32, i32, i32, i32, vp, sz, vp,
                                       vp, vp]
    lib.nf4moe_moe_forward_bias.argtypes = [vp, vp, i32, vp, vp, qp, qp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, sz, vp,
                                            vp, vp]
    lib.nf4moe_ep_pack.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp]
    lib.nf4moe_expert_rows.argtypes = [vp, vp, qp, qp, vp, i32, i32, i32, i32, i32, i32, vp, sz, vp]
    lib.nf4moe_combine_indexed.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
    lib.nf4moe_expert_rows_bias.argtypes = [vp, vp, qp, qp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, sz, vp]
    lib.nf4moe_combine_indexed_bias.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, vp]
    lib.nf4moe_ep_dispatch_p2p.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.nf4moe_ep_return_p2p.argtypes = [vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.nf4moe_ep_wait.argtypes = [vp, i32, vp, vp, vp]
    for name in EXPORTED:
        fn = getattr(lib, name)
        if name not in ("nf4moe_version", "nf4moe_auto_regime", "nf4moe_auto_regime_for", "nf4moe_last_error", "nf4moe_workspace_bytes", "nf4moe_last_launch_count",
                        "nf4moe_workspace_status_offset"):
            fn.restype = i32


def load_library(build_if_missing: bool = True, path: str | None = None):
    """Load (building with nvcc first if the .so is absent or stale) and return the ctypes handle.  `path` loads a differently
    built libnf4moe instead (scripts/ A-B runs of -DNF4MOE_EXPERIMENTS builds); it must be given before the first use -- the
    environment is never consulted."""
    global _lib
    with _lock:
        if _lib is not None:
            if path is not None and os.path.abspath(path) != getattr(_lib, "_nf4moe_path", None):
                raise RuntimeError("libnf4moe is already loaded from another path")
            return _lib
        from . import build as _build
        if path is not None:
            lib = C.CDLL(path)
            _declare(lib)
            lib._nf4moe_path = os.path.abspath(path)
            _lib = lib
            return lib
        if build_if_missing and _build.needs_build():
            try:
                _build.build()
            except Exception as exc:  # stale-but-present library is still usable; absent is fatal
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        "libnf4moe.so is missing and could not be built; this package has no CPU/PyTorch fallback"
                    ) from exc
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libnf4moe.so is missing; run `python gpt-oss-vllm_b200/build.py` (no fallback exists)")
        lib = C.CDLL(LIB_PATH)
        _declare(lib)
        lib._nf4moe_path = os.path.abspath(LIB_PATH)
        _lib = lib
        return lib


def check(rc: int) -> None:
    if rc != 0:
        raise Nf4MoeError(rc, load_library().nf4moe_last_error().decode("utf-8", "replace"))
