"""GPU test (-m gpu) of b2_linear_w8a8_int8_static, the allocation-free form of the fused int8 linear: its scratch
(int8 codes + row scales between the quantizer and the GEMM) is a file-scope `static __device__` array inside the
library (address taken with cudaGetSymbolAddress once per device) instead of a caller workspace.

Called through the C ABI (ctypes) and compared with the existing entry point b2_linear_w8a8_int8 on the same inputs.
The int8 path is bit-exact by contract, so the tolerance is zero differing bits -- the bar tests/test_gpu_parity.py
holds the workspace form to.  Two calls in a row (different shapes, the second smaller, so the shared array still
holds the first call's codes) plus the size limit.

The file name sorts last on purpose: a failure here must not hide other tests.
It passes alone in a fresh process and after the rest of the suite (nothing here depends on earlier state; the array
keeps the last call's contents and nothing in it needs to start at zero).
"""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

import sglang_diffusion_b200 as b2  # noqa: E402

capi = b2.capi
DEV = "cuda:0"


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _problem(M, N, K, seed, dtype=torch.bfloat16, with_bias=True):
    g = torch.Generator(device=DEV).manual_seed(seed)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).to(dtype)
    x[:, ::512] *= 30  # outlier columns: exercises the scale range
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=DEV)
    Bs = torch.rand(N, generator=g, device=DEV) * 1e-2
    bias = torch.randn(N, generator=g, device=DEV).to(dtype) if with_bias else None
    return x, W, Bs, bias


def _dt(t):
    return capi.DTYPE_BF16 if t.dtype == torch.bfloat16 else capi.DTYPE_F16


def linear_workspace(x, W, Bs, bias):
    """The existing entry point: caller-provided workspace."""
    M, K = x.shape
    N = W.shape[0]
    ws = torch.empty(capi.lib.b2_linear_w8a8_int8_workspace(M, K), dtype=torch.uint8, device=DEV)
    D = torch.empty(M, N, dtype=x.dtype, device=DEV)
    capi.check(capi.lib.b2_linear_w8a8_int8(x.data_ptr(), _dt(x), x.stride(0), W.data_ptr(), W.stride(0), Bs.data_ptr(),
                                            None if bias is None else bias.data_ptr(), _dt(x), D.data_ptr(), N, M, N, K,
                                            ws.data_ptr(), ws.numel(), _stream()))
    return D


def linear_static(x, W, Bs, bias, D=None):
    """The new entry point: no workspace argument, nothing allocated inside the call (D is allocated by the caller)."""
    M, K = x.shape
    N = W.shape[0]
    if D is None:
        D = torch.empty(M, N, dtype=x.dtype, device=DEV)
    capi.check(capi.lib.b2_linear_w8a8_int8_static(x.data_ptr(), _dt(x), x.stride(0), W.data_ptr(), W.stride(0),
                                                   Bs.data_ptr(), None if bias is None else bias.data_ptr(), _dt(x),
                                                   D.data_ptr(), N, M, N, K, _stream()))
    return D


def test_static_scratch_linear_equals_workspace_form_on_two_calls_in_a_row():
    # call 1: prefill-sized rows (persistent GEMM), fills 2.1 MB of the array; call 2: decode shape (split-K GEMM),
    # smaller -- the array still holds call 1's codes behind it
    shapes = [(512, 1024, 4096, torch.bfloat16, True), (16, 6144, 4096, torch.bfloat16, False)]
    problems = [_problem(M, N, K, 100 + i, dt, wb) for i, (M, N, K, dt, wb) in enumerate(shapes)]
    want = [linear_workspace(*p) for p in problems]
    outs = [torch.full((p[0].shape[0], p[1].shape[0]), float("nan"), dtype=p[0].dtype, device=DEV) for p in problems]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for p, D in zip(problems, outs):  # back to back on one stream, no synchronize in between
        linear_static(*p, D=D)
    after = torch.cuda.memory_allocated()
    torch.cuda.synchronize()
    assert after == before  # torch's allocator was not touched between the two calls
    for got, ref in zip(outs, want):
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16))
    # and once more in the other order, f16, with the array holding the previous contents
    x, W, Bs, bias = _problem(100, 520, 1168, 7, torch.float16, True)
    assert torch.equal(linear_static(x, W, Bs, bias).view(torch.int16), linear_workspace(x, W, Bs, bias).view(torch.int16))


def test_static_scratch_linear_matches_torch_op_at_the_size_limit():
    limit = capi.lib.b2_linear_w8a8_int8_static_max_bytes()
    assert limit == 4 << 20
    M, K, N = 128, 28672, 256  # 3.67 MB of codes: the largest decode-class activation of BASELINE C2
    assert capi.lib.b2_linear_w8a8_int8_workspace(M, K) <= limit
    x, W, Bs, bias = _problem(M, N, K, 11)
    got = linear_static(x, W, Bs, bias)
    ref = b2.int8_scaled_mm_with_quant(x, W, Bs, bias, torch.bfloat16, False)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), ref.view(torch.int16))


def test_static_scratch_linear_rejects_sizes_beyond_the_array():
    M, K, N = 1024, 4096, 256  # 4 MiB of codes + scales: just over
    assert capi.lib.b2_linear_w8a8_int8_workspace(M, K) > capi.lib.b2_linear_w8a8_int8_static_max_bytes()
    x, W, Bs, bias = _problem(M, N, K, 3)
    D = torch.full((M, N), 7.0, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(capi.B2Error) as ei:
        linear_static(x, W, Bs, bias, D=D)
    assert "B2_STATIC_SCRATCH_BYTES" in str(ei.value) and ei.value.code == capi.ERR_ARG
    torch.cuda.synchronize()
    assert bool((D == 7.0).all())  # nothing was launched
    # the library stays usable
    x, W, Bs, bias = _problem(8, 256, 512, 4)
    assert torch.equal(linear_static(x, W, Bs, bias).view(torch.int16), linear_workspace(x, W, Bs, bias).view(torch.int16))
