"""Host-side mirror of the reference's quant-method interface for the NF4 fused-MoE expert path.

* `B200NF4MoEConfig`   <- P/bnb_moe_4bit.py:22-56 (`BitsAndBytesMoE4bitConfig`): same name/fields/targeting rule, but
                          `get_quant_method` returns a real method object instead of the string "bnb_4bit".
* `B200NF4MoEMethod`   <- vLLM `BitsAndBytesMoEMethod` (V/model_executor/layers/quantization/bitsandbytes.py:435-603):
                          `create_weights` registers the same parameters; `apply` (0.22 split signature),
                          `apply_monolithic` and the legacy `(x, router_logits, top_k, renormalize)` form all call the
                          same CUDA kernels.
* `NF4MoELayer`        <- P/unsloth_to_vllm_converter.py:270-325 (`UnslothMoEWrapper`): `forward(hidden_states,
                          router_logits, top_k)` with [B,S,H] inputs.

vLLM is not imported here; `vllm_plugin.py` does the registration when vLLM is present.
"""
from __future__ import annotations

import torch

from . import _lib, ops
from .quant import NF4ExpertWeights

ACTIVATIONS = {"swigluoai": _lib.ACT_SWIGLU_OAI, "swiglu_oai": _lib.ACT_SWIGLU_OAI, "chunk_silu": _lib.ACT_CHUNK_SILU,
               "silu": _lib.ACT_CHUNK_SILU}


class B200NF4MoEConfig:
    """Mirror of P/bnb_moe_4bit.py:22-56.  blocksize 64 is authoritative (the reference's group_size=128 at
    :27 contradicts every loader it ships)."""

    def __init__(self, weight_bits: int = 4, group_size: int = 64, compute_dtype: torch.dtype = torch.float16,
                 use_nf4: bool = True):
        if weight_bits != 4 or not use_nf4 or group_size != 64:
            raise ValueError("this path is NF4, 4 bits, blocksize 64")
        self.weight_bits, self.group_size, self.compute_dtype, self.use_nf4 = weight_bits, group_size, compute_dtype, use_nf4

    def get_name(self) -> str:  # P/bnb_moe_4bit.py:33
        return "bnb_moe_4bit"

    def get_supported_act_dtypes(self):  # :37 lists fp16/bf16; the kernels take fp16 operands, bf16 is widened exactly (ops.moe_forward)
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:  # :54-56 says 70; this build is sm_100a only
        return 100

    @staticmethod
    def is_moe_weight(prefix: str) -> bool:  # P/bnb_moe_4bit.py:43-48, 99-118
        return "experts" in prefix and any(n in prefix for n in ("w13", "w2", "gate", "up", "down"))

    def get_quant_method(self, layer, prefix: str):
        return B200NF4MoEMethod(self) if self.is_moe_weight(prefix) or prefix.endswith("experts") else None


class B200NF4MoEMethod:
    """FusedMoE quant method backed by libnf4moe."""

    is_monolithic = True  # routing can run inside (apply_monolithic); `apply` accepts precomputed routing too

    def __init__(self, quant_config: B200NF4MoEConfig | None = None, moe=None, top_k: int | None = None,
                 activation: str = "swigluoai", nested: bool | None = None, regime: int = _lib.REGIME_AUTO):
        self.quant_config = quant_config or B200NF4MoEConfig()
        self.moe = moe
        self.top_k = top_k
        self.act = ACTIVATIONS[activation]
        self.nested = nested
        self.regime = regime
        _lib.load_library()  # fail at construction, not at first forward, if the CUDA library is missing

    # -- V/.../quantization/bitsandbytes.py:451-471, 508-575: identical parameter names, shapes and attributes --
    def create_weights(self, layer: torch.nn.Module, num_experts: int, hidden_size: int,
                       intermediate_size_per_partition: int, params_dtype: torch.dtype, has_bias: bool = False,
                       **extra_weight_attrs):
        quant_ratio = 2  # two 4-bit codes per uint8
        I = intermediate_size_per_partition
        if has_bias:
            # gpt-oss expert biases under vLLM's names (V/model_executor/layers/fused_moe/unquantized_fused_moe_method.py
            # registers the same two for the unquantised path; vLLM's bnb method has none)
            for name, width in (("w13_bias", 2 * I), ("w2_bias", hidden_size)):
                b = torch.nn.Parameter(torch.zeros(num_experts, width, dtype=params_dtype), requires_grad=False)
                for k, v in extra_weight_attrs.items():
                    setattr(b, k, v)
                layer.register_parameter(name, b)
        for name, out_dim, in_dim in (("w13_weight", 2 * I, hidden_size), ("w2_weight", hidden_size, I)):
            p = torch.nn.Parameter(torch.empty(num_experts, out_dim * in_dim // quant_ratio, 1, dtype=torch.uint8),
                                   requires_grad=False)
            attrs = dict(extra_weight_attrs)
            attrs.update(num_experts=num_experts, input_dim=in_dim, output_dim=out_dim,
                         experts_shape=(num_experts, out_dim, in_dim), pack_factor=quant_ratio,
                         use_bitsandbytes_4bit=True)
            for k, v in attrs.items():
                setattr(p, k, v)
            layer.register_parameter(name, p)

    def bind(self, layer: torch.nn.Module, w13: NF4ExpertWeights, w2: NF4ExpertWeights) -> None:
        """Attach already-packed weights (no re-quantisation, no unpacking: P/unsloth_bnb_analysis.md:57)."""
        layer._nf4_w13, layer._nf4_w2 = w13, w2

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        if not hasattr(layer, "_nf4_w13"):
            self.bind(layer, NF4ExpertWeights.from_vllm_param(layer.w13_weight),
                      NF4ExpertWeights.from_vllm_param(layer.w2_weight))
        layer._nf4_bias = None   # (re)built from the layer's bias parameters on the next call

    def _weights(self, layer):
        if not hasattr(layer, "_nf4_w13"):
            self.process_weights_after_loading(layer)
        return layer._nf4_w13, layer._nf4_w2

    @staticmethod
    def _biases(layer) -> dict:
        """gpt-oss expert biases under vLLM's parameter names (`w13_bias [E, 2I]`, `w2_bias [E, H]`,
        V/model_executor/models/gpt_oss.py:199), when the layer carries them.  The fp16 contiguous form the
        kernels read is made ONCE per source tensor and cached on the layer (a cast per call would allocate on the hot path and
        break CUDA-graph capture); replacing the parameter's storage (a reload) invalidates the cache."""
        cache = getattr(layer, "_nf4_bias", None)
        out, key = {}, []
        for name in ("w13_bias", "w2_bias"):
            b = getattr(layer, name, None)
            if b is not None:
                b = b.data if isinstance(b, torch.nn.Parameter) else b
                key.append((name, b.data_ptr(), b.dtype, tuple(b.shape)))
        key = tuple(key)
        if cache is not None and cache[0] == key:
            return cache[1]
        for name in ("w13_bias", "w2_bias"):
            b = getattr(layer, name, None)
            if b is not None:
                b = b.data if isinstance(b, torch.nn.Parameter) else b
                out[name] = b if (b.dtype == torch.float16 and b.is_contiguous()) else b.to(torch.float16).contiguous()
        try:
            layer._nf4_bias = (key, out)
        except Exception:   # a layer object that refuses new attributes: recompute per call
            pass
        return out

    def _act(self, layer) -> int:
        a = getattr(layer, "activation", None)
        if a is None:
            return self.act
        return ACTIVATIONS[getattr(a, "value", a)]

    # -- vLLM 0.22 split signature, V/.../fused_moe_method_base.py:163-172 --
    def apply(self, layer, x: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor, shared_experts=None,
              shared_experts_input=None) -> torch.Tensor:
        """Routing pre-computed.  Honours the pass-through arguments vLLM's bnb method forwards to `fused_experts`
        (V/model_executor/layers/quantization/bitsandbytes.py:493-505), read from the layer as vLLM does:
          * `expert_map` / `global_num_experts` (vLLM's own expert parallelism): global ids are mapped to this rank's local
            ids, foreign experts become -1 and are skipped by the bucket and combine kernels;
          * `apply_router_weight_on_input`: the routing weight multiplies the token BEFORE the experts (top_k must be 1, as in
            vLLM's fused_experts) and the combine then sums with weight 1."""
        w13, w2 = self._weights(layer)
        expert_map = getattr(layer, "expert_map", None)
        if expert_map is not None:
            gne = getattr(layer, "global_num_experts", None)
            if gne is not None and gne > 0 and expert_map.numel() != gne:
                raise ValueError(f"expert_map has {expert_map.numel()} entries for global_num_experts={gne}")
            topk_ids = expert_map[topk_ids.long()].to(torch.int32)
        if getattr(layer, "apply_router_weight_on_input", False):
            if topk_ids.shape[1] != 1:
                raise ValueError("apply_router_weight_on_input is only defined for top_k == 1 (as in vLLM's fused_experts)")
            x = (x * topk_weights.to(x.dtype)).contiguous()
            topk_weights = torch.ones_like(topk_weights, dtype=torch.float32)
        return ops.moe_forward(x, w13, w2, topk_ids.shape[1], topk_ids=topk_ids, topk_weights=topk_weights,
                               act=self._act(layer), regime=self.regime, nested=self.nested, **self._biases(layer))

    # -- V/.../fused_moe_method_base.py:174-181 --
    def apply_monolithic(self, layer, x: torch.Tensor, router_logits: torch.Tensor, input_ids=None) -> torch.Tensor:
        top_k = self.top_k or getattr(layer, "top_k", None) or getattr(self.moe, "experts_per_token", None)
        if top_k is None:
            raise ValueError("top_k unknown: pass top_k= to the method or set layer.top_k")
        return self.apply_legacy(layer, x, router_logits, top_k, renormalize=True)

    # -- the north-star / Aug-2025 `FusedMoEMethodBase.apply(layer, x, router_logits, top_k, renormalize)` shape --
    def apply_legacy(self, layer, x: torch.Tensor, router_logits: torch.Tensor, top_k: int,
                     renormalize: bool = True) -> torch.Tensor:
        if not renormalize:
            raise NotImplementedError("gpt-oss routes with renormalize=True (V/model_executor/models/gpt_oss.py:195)")
        w13, w2 = self._weights(layer)
        return ops.moe_forward(x, w13, w2, top_k, router_logits=router_logits.contiguous(), act=self._act(layer),
                               regime=self.regime, nested=self.nested, **self._biases(layer))


class NF4MoELayer(torch.nn.Module):
    """Stand-alone layer with the reference wrapper's call shape (P/unsloth_to_vllm_converter.py:282)."""

    def __init__(self, w13: NF4ExpertWeights, w2: NF4ExpertWeights, top_k: int = 4, activation: str = "swigluoai",
                 nested: bool | None = None, regime: int = _lib.REGIME_AUTO):
        super().__init__()
        self.num_experts = w13.num_experts
        self.hidden_size, self.intermediate_size = w13.cols, w13.rows // 2
        self.top_k = top_k
        self.method = B200NF4MoEMethod(top_k=top_k, activation=activation, nested=nested, regime=regime)
        self.method.bind(self, w13, w2)

    def forward(self, hidden_states: torch.Tensor, router_logits: torch.Tensor, top_k: int | None = None):
        shape = hidden_states.shape
        x = hidden_states.reshape(-1, shape[-1]).contiguous()
        lg = router_logits.reshape(-1, router_logits.shape[-1]).contiguous()
        out = self.method.apply_legacy(self, x, lg, top_k or self.top_k)
        return out.reshape(shape)
