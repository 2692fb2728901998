"""nf4moe_quantize_static: the NF4 quantiser for callers that cannot allocate device memory.

Who needs it: a weight loader inside a serving process whose HBM an engine has already handed to its KV-cache pool (a cudaMalloc
for the quantiser's temporaries may then fail although the layer fits), and C / ctypes hosts of the ABI that have no allocator of
their own or may not call cudaMalloc / cudaFree while requests are in flight, because both can synchronise the device.  The
entry point keeps the quantiser's scratch in a fixed 4 MiB array inside the library instead of taking a caller buffer.

What is checked, through the C ABI only: for two calls in a row (different shapes, so the second call finds the first call's
leftovers in the shared array) every output equals nf4moe_quantize's bit for bit -- the tolerance the quantiser's other tests
use; the second call leaves the device's free memory where it was; a request beyond the array's size is refused with both sizes
in the message and writes nothing.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the parity tests.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 4194304   # NF4MOE_STATIC_SCRATCH_BYTES in include/nf4moe.h


def _code2():
    # any sorted 256-entry table serves: both entry points read the same one
    return torch.linspace(-1.0, 1.0, 256, dtype=torch.float32)


def _outputs(E, rows, cols, dev):
    nb = rows * cols // 64
    n2 = -(-nb // 256)
    return dict(packed=torch.full((E, rows * cols // 2), 0xAB, dtype=torch.uint8, device=dev),
                absmax_u8=torch.full((E, nb), 0xCD, dtype=torch.uint8, device=dev),
                absmax2=torch.full((E, n2), -7.0, dtype=torch.float32, device=dev),
                offset=torch.full((E,), -7.0, dtype=torch.float32, device=dev),
                absmax=torch.full((E, nb), -7.0, dtype=torch.float32, device=dev))


def _call(lib, check, static, w, code2, out, stream):
    E, rows, cols = w.shape
    args = [w.data_ptr(), E, rows, cols, out["packed"].data_ptr(), out["absmax_u8"].data_ptr(), out["absmax2"].data_ptr(),
            code2.data_ptr(), out["offset"].data_ptr(), out["absmax"].data_ptr()]
    if static:
        return lib.nf4moe_quantize_static(*args, stream)
    nb = rows * cols // 64
    scratch = torch.empty(E * nb * 4 + E * 8 + 512, dtype=torch.uint8, device=w.device)
    rc = lib.nf4moe_quantize(*args, scratch.data_ptr(), stream)
    torch.cuda.synchronize(w.device)      # `scratch` must outlive the two passes
    return rc


def test_quantize_static_matches_quantize_on_two_calls_in_a_row(pkg, cuda_device):
    lib = pkg._lib.load_library()
    check = pkg._lib.check
    dev = cuda_device
    g = torch.Generator().manual_seed(20250805)
    code2 = _code2().to(dev)
    # first call: 3 experts with a partial last nested block and the quantiser's edge cases; second call: a different, larger
    # geometry (one full-size gpt-oss down matrix), so every scratch word it reads was last written by the first call or never
    cases = [(3, 160, 448), (1, 2880, 2880)]
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        for call, (E, rows, cols) in enumerate(cases):
            w = (torch.randn(E, rows, cols, generator=g) * 0.02).half()
            w[0, 0, :64] = 0            # all-zero block
            w[E - 1, 3, 70] = 2.5       # one outlier in a block
            w = w.to(dev)
            want, got = _outputs(E, rows, cols, dev), _outputs(E, rows, cols, dev)
            check(_call(lib, check, False, w, code2, want, stream))
            torch.cuda.synchronize(dev)
            free_before = torch.cuda.mem_get_info(dev)[0]
            check(_call(lib, check, True, w, code2, got, stream))
            torch.cuda.synchronize(dev)
            free_after = torch.cuda.mem_get_info(dev)[0]
            if call > 0:                # the first call may load the module and look the array up; later ones touch nothing
                assert free_after >= free_before, "nf4moe_quantize_static took device memory"
            for name in want:
                a, b = want[name].cpu().numpy(), got[name].cpu().numpy()
                assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"call {call}: {name} differs from nf4moe_quantize"
            assert not torch.equal(got["absmax"], torch.full_like(got["absmax"], -7.0))   # the outputs were really written


def test_quantize_static_refuses_what_its_array_cannot_hold(pkg, cuda_device):
    lib = pkg._lib.load_library()
    dev = cuda_device
    E, rows, cols = 5, 5760, 2880        # 5 gate/up matrices: 5 * 259,200 absmax words + sums > 4 MiB (4 of them fit)
    need = 256 + E * (rows * cols // 64) * 4
    assert need > STATIC_SCRATCH_BYTES >= 256 + 4 * (rows * cols // 64) * 4
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        w = torch.zeros(8, dtype=torch.float16, device=dev)        # never read: the size check comes first
        out = _outputs(1, 64, 64, dev)
        before = {k: v.clone() for k, v in out.items()}
        code2 = _code2().to(dev)
        rc = lib.nf4moe_quantize_static(w.data_ptr(), E, rows, cols, out["packed"].data_ptr(), out["absmax_u8"].data_ptr(),
                                        out["absmax2"].data_ptr(), code2.data_ptr(), out["offset"].data_ptr(),
                                        out["absmax"].data_ptr(), stream)
        torch.cuda.synchronize(dev)
    assert rc == -1                       # NF4MOE_EINVAL
    msg = lib.nf4moe_last_error().decode()
    assert str(need) in msg and str(STATIC_SCRATCH_BYTES) in msg, msg
    for k in out:
        assert torch.equal(out[k], before[k])
    header = open(os.path.join(ROOT, "include", "nf4moe.h")).read()
    assert f"#define NF4MOE_STATIC_SCRATCH_BYTES {STATIC_SCRATCH_BYTES}" in header
