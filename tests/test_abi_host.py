"""CPU: the C-ABI library loads and exports every symbol include/nf4moe.h declares; host-side logic; the product
path never touches the oracle and fails loudly without a GPU (no compute calls are made here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols(pkg):
    lib = pkg.load_library()
    header = open(os.path.join(ROOT, "include", "nf4moe.h")).read()
    body = header[header.index("typedef struct"):]                       # past the comment block that lists the ops
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    declared = set(re.findall(r"\b(nf4moe_[a-z0-9_]+)\s*\(", body))
    assert declared == set(pkg._lib.EXPORTED)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.nf4moe_version() == 101 == pkg._lib.ABI_VERSION
    assert int(re.search(r"#define NF4MOE_VERSION (\d+)", header).group(1)) == 101


def test_qweight_struct_layout_matches_header(pkg):
    q = pkg._lib.QWeight
    assert [f[0] for f in q._fields_] == ["packed", "absmax_f32", "absmax_u8", "absmax2", "code2", "offset",
                                          "code2_stride", "rows", "cols"]
    assert C.sizeof(q) == 6 * 8 + 8 + 4 + 4


def test_usage_errors_return_codes_without_gpu(pkg):
    lib = pkg.load_library()
    q = pkg._lib.QWeight()
    assert lib.nf4moe_dequant(C.byref(q), 0, 1, None, None) == -1          # NF4MOE_EINVAL: null packed
    assert b"null" in lib.nf4moe_last_error()
    q.packed, q.rows, q.cols = 16, 4, 100                                   # cols not a multiple of 64
    assert lib.nf4moe_dequant(C.byref(q), 0, 1, 16, None) == -1
    q.cols = 64                                                             # no absmax form
    assert lib.nf4moe_dequant(C.byref(q), 0, 1, 16, None) == -1
    q.absmax_f32, q.packed = 16, 8                                          # misaligned packed
    assert lib.nf4moe_dequant(C.byref(q), 0, 1, 16, None) == -2            # NF4MOE_EALIGN
    assert lib.nf4moe_route_topk(None, 0, 4, 32, 4, None, None, None) == -1
    assert lib.nf4moe_moe_forward(None, None, 0, None, None, None, None, None, 4, 32, 64, 64, 4, 0, 0, None, 0, None,
                                  None, None) == -1


def test_workspace_bytes(pkg):
    ws = pkg.ops.workspace_bytes
    assert ws(0, 32, 2880, 2880, 4) > 0 and ws(64, 32, 2880, 2880, 4) > ws(1, 32, 2880, 2880, 4)
    assert ws(8192, 32, 2880, 2880, 4) > 8192 * 4 * 2880 * 2 * 3
    assert ws(1, 0, 1, 1, 1) == 0


def test_status_word_and_pool_contract(pkg):
    lib = pkg.load_library()
    assert lib.nf4moe_workspace_status_offset(64, 32, 2880, 2880, 4) == 0      # the first word, for every geometry
    assert lib.nf4moe_workspace_status_offset(8192, 128, 2880, 2880, 4) == 0
    pool = pkg.ops.WorkspacePool()
    assert pool.buffers() == []
    assert (pkg._lib.STATUS_EP_OVERFLOW, pkg._lib.STATUS_DEC_NOFIT, pkg._lib.STATUS_GEMM_NOFIT) == (1, 2, 4)
    header = open(os.path.join(ROOT, "include", "nf4moe.h")).read()
    assert (pkg._lib.STATUS_EP_TIMEOUT, pkg._lib.STATUS_DEP_TIMEOUT) == (8, 16)
    for name, val in (("NF4MOE_STATUS_EP_OVERFLOW", 1), ("NF4MOE_STATUS_DEC_NOFIT", 2), ("NF4MOE_STATUS_GEMM_NOFIT", 4),
                      ("NF4MOE_STATUS_EP_TIMEOUT", 8), ("NF4MOE_STATUS_DEP_TIMEOUT", 16)):
        assert re.search(rf"#define {name} {val}\b", header)


def test_shipped_library_reads_no_environment_on_the_launch_path():
    """Experiment switches (ring depths, tile sizes, kernel variants) exist only behind -DNF4MOE_EXPERIMENTS: a stray
    environment variable must never change the numerics or the kernel choice of the drop-in library."""
    csrc = os.path.join(ROOT, "gpt-oss-vllm_b200", "csrc")
    for f in os.listdir(csrc):
        src = open(os.path.join(csrc, f)).read()
        # every getenv sits inside an #ifdef NF4MOE_EXPERIMENTS block
        depth, guarded = 0, True
        stack = []
        for line in src.splitlines():
            t = line.strip()
            if t.startswith("#if"):
                stack.append("NF4MOE_EXPERIMENTS" in t and not t.startswith("#ifndef"))
            elif t.startswith("#else") and stack:
                stack[-1] = False
            elif t.startswith("#endif") and stack:
                stack.pop()
            elif "getenv(" in t and not t.startswith("//"):
                assert any(stack), f"{f}: getenv outside NF4MOE_EXPERIMENTS: {t}"
    assert "NF4MOE_LIB" not in open(os.path.join(ROOT, "gpt-oss-vllm_b200", "_lib.py")).read()


def test_no_global_load_ahead_of_the_grid_dependency_wait(pkg):
    """Kernels launched with programmatic dependent launch start while their predecessor is still running; everything they
    read from it must be loaded AFTER griddepcontrol.wait (SASS: ACQBULK).  The compiler may hoist loads through
    `const __restrict__` pointers above the wait (round 2: permute_rows_kernel read stale bucket metadata that way), so the built
    library's SASS is scanned: no global / TMA load may precede the first ACQBULK, except the decode kernel's preload of the
    constant code2 table (weights, written before any launch)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        # the source-level rule (common.cuh: no __restrict__ / __ldg on anything an earlier kernel writes) is the first line of
        # defence; this scan is the second, and it must not vanish quietly where the kernels actually run
        assert not torch.cuda.is_available(), "cuobjdump is missing on a GPU box: the PDL load-order scan cannot run"
        pytest.skip("cuobjdump not available (no GPU here either)")
    lib = os.path.join(os.path.dirname(pkg.__file__), "libnf4moe.so")
    sass = subprocess.run([cuobjdump, "-sass", lib], capture_output=True, text=True, check=True).stdout
    checked = 0
    for part in re.split(r"\n\s*Function : ", sass)[1:]:
        name = part.split("\n", 1)[0]
        ins = re.findall(r"/\*[0-9a-f]{4,}\*/\s+(.*?);", part)
        acq = [i for i, t in enumerate(ins) if "ACQBULK" in t]
        if not acq:
            continue
        checked += 1
        early = [t.strip() for t in ins[:acq[0]] if re.search(r"\bLDG|\bLD\.|UTMALDG|UBLKCP|LDGSTS", t)]
        # the decode kernel's FAST variants (second template argument true) preload the constant code2 table ahead of the wait:
        # ceil(256 * 32 / 640 threads) = 13 loads per thread, all of weights; nothing else may sit there
        allowed = 13 if re.search(r"dec_kernelILi\dELb1E", name) else 0
        assert len(early) <= allowed, f"{name}: loads ahead of the dependency wait: {early}"
    assert checked >= 10, "expected the PDL kernels (route, bucket, permute, GEMV, tcgen05, combine) in the library"


def test_no_restrict_or_ldg_on_buffers_an_earlier_kernel_writes():
    """Source-level form of the same rule: in the kernels' parameter lists, `const ... __restrict__` may only name weights, absmax,
    code tables and biases, and __ldg only reads those; everything a preceding kernel may have produced is a plain pointer."""
    csrc = os.path.join(ROOT, "gpt-oss-vllm_b200", "csrc")
    produced = r"\b(x|xs|y|y_rows|h|raw|logits|ids|recv_ids|w|weights|expert_offsets|sorted_pairs|active_experts|meta|tile_list|pair_slot|src)\b"
    for f in sorted(os.listdir(csrc)):
        text = open(os.path.join(csrc, f)).read()
        for m in re.finditer(r"const\s+[\w:]+\s*\*\s*__restrict__\s+(\w+)", text):
            name = m.group(1)
            if f == "dequant.cuh" and name in ("w", "src"):     # the quantiser / cast kernels: `w` is the fp16 weight input of plain
                line = text[:m.start()].count("\n") + 1         # (non-PDL) launches; the cast kernel reads src through ld_dep_u4
                assert "quantize_blocks_kernel" in text.splitlines()[line - 1] or "cast_f16_bf16_kernel" in text.splitlines()[line - 1]
                continue
            assert not re.fullmatch(produced, name), f"{f}: `{name}` is written by earlier kernels but is const __restrict__"
        for m in re.finditer(r"__ldg\(([^;]*);", text):
            arg = m.group(1)
            assert re.search(r"code2|absmax|offset|bias|\bbe\b", arg), f"{f}: __ldg on something that is not a weight-side table: {arg}"


def test_config_and_create_weights_mirror_reference(pkg):
    cfg = pkg.B200NF4MoEConfig()
    assert cfg.get_name() == "bnb_moe_4bit"                                 # P/bnb_moe_4bit.py:33
    assert cfg.is_moe_weight("model.layers.3.mlp.experts.w13_weight")       # P/bnb_moe_4bit.py:43-48
    assert cfg.is_moe_weight("model.layers.3.mlp.experts.down_proj")
    assert not cfg.is_moe_weight("model.layers.3.self_attn.qkv_proj")
    assert cfg.get_quant_method(None, "model.layers.0.self_attn.o_proj") is None
    m = cfg.get_quant_method(None, "model.layers.0.mlp.experts")
    layer = torch.nn.Module()
    m.create_weights(layer, 32, 2880, 2880, torch.float16, weight_loader="wl")
    # P/validate_conversion.py:45-46, P/debug_shapes.py:53-54
    assert tuple(layer.w13_weight.shape) == (32, 8294400, 1) and tuple(layer.w2_weight.shape) == (32, 4147200, 1)
    assert layer.w13_weight.experts_shape == (32, 5760, 2880) and layer.w2_weight.experts_shape == (32, 2880, 2880)
    assert layer.w13_weight.pack_factor == 2 and layer.w13_weight.use_bitsandbytes_4bit and layer.w13_weight.weight_loader == "wl"
    n = sum(int(np.prod(p.experts_shape)) for p in (layer.w13_weight, layer.w2_weight))
    assert n == 32 * (5760 * 2880 + 2880 * 2880)                            # P/validate_conversion.py:187
    with pytest.raises(ValueError):
        pkg.B200NF4MoEConfig(group_size=128)


def test_expert_weights_validation_and_bytes(pkg):
    packed = torch.zeros(2, 64 * 128 // 2, 1, dtype=torch.uint8)
    w = pkg.NF4ExpertWeights(packed, 64, 128, absmax=torch.zeros(2, 128))
    assert w.num_experts == 2 and w.blocks_per_expert == 128 and not w.has_nested
    assert w.nbytes(nested=False) == 4096 + 512 and w.nbytes(nested=True) == 4096 + 128 + 4
    with pytest.raises(ValueError):
        pkg.NF4ExpertWeights(packed, 64, 100, absmax=torch.zeros(2, 100))
    with pytest.raises(ValueError):
        pkg.NF4ExpertWeights(packed, 64, 128)
    with pytest.raises(ValueError):
        w.c_struct(nested=True)
    # 20b sizes (SURVEY 8 header)
    big = pkg.NF4ExpertWeights.__new__(pkg.NF4ExpertWeights)
    big.rows, big.cols = 5760, 2880
    assert big.blocks_per_expert == 259200 and big.nbytes(False) == 8294400 + 1036800
    assert big.nbytes(True) == 8294400 + 259200 + 4 * 1013


def test_no_cpu_fallback(pkg):
    packed = torch.zeros(1, 64 * 64 // 2, 1, dtype=torch.uint8)
    w = pkg.NF4ExpertWeights(packed, 64, 64, absmax=torch.ones(1, 64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        w.dequantize()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.route_topk(torch.zeros(2, 32, dtype=torch.float16), 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.ops.moe_forward(torch.zeros(2, 64, dtype=torch.float16), w, w, 4,
                            router_logits=torch.zeros(2, 32, dtype=torch.float16))


def test_product_never_imports_oracle():
    pk = os.path.join(ROOT, "gpt-oss-vllm_b200")
    for dirpath, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "nf4_oracle" not in src and not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f


def test_alu_decode_tables_are_the_fp16_nf4_codes():
    """csrc/gemm_tc.cuh decode_word_alu (the -DNF4MOE_DEC_ARITH path): its PRMT byte-plane constants and selectors, emulated on
    the host, give the same half2 per packed byte as the look-up table (element 2j = HIGH nibble in the low half)."""
    import re
    import numpy as np
    from oracle import nf4_oracle as orc
    src = open(os.path.join(ROOT, "gpt-oss-vllm_b200", "csrc", "gemm_tc.cuh")).read()
    c = {m.group(1): int(m.group(2), 16) for m in re.finditer(r"(kNF4(?:Hi|Lo)[A-D]) = (0x[0-9a-f]+)u", src)}
    assert len(c) == 8
    h16 = np.asarray(orc.NF4_CODE, np.float32).astype(np.float16).view(np.uint16)

    def prmt(a, b, sel):
        by = [(a >> (8 * i)) & 255 for i in range(4)] + [(b >> (8 * i)) & 255 for i in range(4)]
        r = 0
        for i in range(4):
            s = (sel >> (4 * i)) & 15
            v = by[s & 7]
            if s & 8:
                v = 0xFF if v & 0x80 else 0
            r |= v << (8 * i)
        return r

    def decode(w):
        s, w4, out = w & 0x77777777, (w << 4) & 0xFFFFFFFF, []
        for msel, sel in ((0xD9C8, s), (0xFBEA, s >> 16)):
            m = prmt(w4, w, msel)
            rh = (prmt(c["kNF4HiA"], c["kNF4HiB"], sel) & ~m & 0xFFFFFFFF) | (prmt(c["kNF4HiC"], c["kNF4HiD"], sel) & m)
            rl = (prmt(c["kNF4LoA"], c["kNF4LoB"], sel) & ~m & 0xFFFFFFFF) | (prmt(c["kNF4LoC"], c["kNF4LoD"], sel) & m)
            out += [prmt(rl, rh, 0x4051), prmt(rl, rh, 0x6273)]
        return out

    rng = np.random.default_rng(3)
    for w in [0, 0xFFFFFFFF, 0x01234567, 0x89ABCDEF, 0xFEDCBA98, 0x76543210] + [int(v) for v in rng.integers(0, 2 ** 32, 4000)]:
        ref = [int(h16[((w >> (8 * j)) & 255) >> 4]) | (int(h16[(w >> (8 * j)) & 15]) << 16) for j in range(4)]
        assert decode(w) == ref, hex(w)
