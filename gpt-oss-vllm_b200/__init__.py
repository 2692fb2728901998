"""B200-native NF4 fused-MoE expert path (drop-in for the bitsandbytes FusedMoE quant method of
toddllm/gpt-oss-vllm).  Host code is a thin Python mirror of the reference interface; all arithmetic runs in
hand-written sm_100a CUDA kernels behind the C ABI of `include/nf4moe.h` (libnf4moe.so).  No CPU fallback.

The directory name has hyphens (it follows the reference's name), so import it with
    importlib.import_module("gpt-oss-vllm_b200")        # repo root on sys.path
or through the alias module `nf4moe_b200` at the repo root.
"""
from . import _lib, ops  # noqa: F401
from ._lib import (ACT_CHUNK_SILU, ACT_SWIGLU_OAI, REGIME_AUTO, REGIME_GEMM, REGIME_GEMM_FAST, REGIME_GEMV, Nf4MoeError,  # noqa: F401
                   load_library)
from .graphed import GraphedForward  # noqa: F401
from .moe_method import B200NF4MoEConfig, B200NF4MoEMethod, NF4MoELayer  # noqa: F401
from .quant import BLOCKSIZE, NESTED_BLOCKSIZE, NF4ExpertWeights, dynamic_code_map  # noqa: F401

__all__ = ["NF4ExpertWeights", "B200NF4MoEConfig", "B200NF4MoEMethod", "NF4MoELayer", "GraphedForward", "ops", "load_library",
           "Nf4MoeError"]
