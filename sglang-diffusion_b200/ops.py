"""Python wrappers with the names and signatures of the reference's wrappers.

  sgl-kernel/python/sgl_kernel/gemm.py:13-42,154-159   int8_scaled_mm, fp8_scaled_mm,
                                                       fp8_blockwise_scaled_mm, sgl_per_token_quant_fp8
  python/sglang/srt/layers/quantization/int8_kernel.py:59-89     per_token_quant_int8
  python/sglang/srt/layers/quantization/fp8_kernel.py:435-541,587-603
                                                       sglang_per_token_group_quant_fp8, sglang_per_token_quant_fp8

Every function is a thin call of torch.ops.sgl_kernel.* (registered by b200_w8a8_torch.so);
nothing here computes.  Fake (meta) kernels are registered below the way the reference's
callers register them (w8a8_int8.py:46-57, fp8_utils.py:69-81) so the ops trace under
torch.compile without sglang being imported.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

fp8_dtype = torch.float8_e4m3fn
fp8_max = torch.finfo(fp8_dtype).max
fp8_min = -fp8_max

_ops = torch.ops.sgl_kernel


# ------------------------------------------------------------------ GEMMs ----

def int8_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None):
    return _ops.int8_scaled_mm.default(mat_a, mat_b, scales_a, scales_b, out_dtype, bias)


def fp8_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None):
    return _ops.fp8_scaled_mm.default(mat_a, mat_b, scales_a, scales_b, out_dtype, bias)


def fp8_blockwise_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype):
    return _ops.fp8_blockwise_scaled_mm.default(mat_a, mat_b, scales_a, scales_b, out_dtype)


def scaled_mm_out(mat_a, mat_b, scales_a, scales_b, out, bias=None):
    """int8_scaled_mm / fp8_scaled_mm (chosen by mat_a.dtype) writing into the caller's `out` [M,N] instead of a
    fresh tensor -- the C ABI takes the output pointer (include/b200_w8a8.h); the torch ops above cannot, their
    schema is the reference's.  Used to place row-parallel partial outputs in the symmetric all-reduce buffer."""
    from . import capi

    M, K = mat_a.shape
    N = mat_b.shape[1]
    assert mat_b.shape[0] == K and mat_b.stride(0) == 1, "mat_b must be [K,N] with stride(0) == 1"
    assert out.shape == (M, N) and out.stride(1) == 1 and mat_a.stride(1) == 1
    assert scales_a.numel() == M and scales_b.numel() == N and scales_a.is_contiguous() and scales_b.is_contiguous()
    dt = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}
    fn = capi.lib.b2_gemm_s8 if mat_a.dtype == torch.int8 else capi.lib.b2_gemm_e4m3
    with torch.cuda.device(mat_a.device):  # the library plans for and launches on the CURRENT device
        stream = torch.cuda.current_stream(mat_a.device).cuda_stream
        capi.check(fn(mat_a.data_ptr(), mat_a.stride(0), mat_b.data_ptr(), mat_b.stride(1), scales_a.data_ptr(),
                      scales_b.data_ptr(), None if bias is None else bias.data_ptr(),
                      dt[bias.dtype] if bias is not None else capi.DTYPE_F32, out.data_ptr(), out.stride(0),
                      dt[out.dtype], M, N, K, stream))
    return out


def fp8_blockwise_scaled_mm_bias(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None):
    """fp8_blockwise_scaled_mm with the bias add fused into the GEMM epilogue.  The reference op has no bias
    parameter (its caller adds it with a second kernel, fp8_utils.py:546-547); the C ABI has one, so the linear
    method calls b2_gemm_e4m3_block128 directly when a bias is present.  Same argument checks as the torch op."""
    from . import capi

    if bias is None:
        return fp8_blockwise_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype)
    M, K = mat_a.shape
    N = mat_b.shape[1]
    assert mat_b.shape[0] == K and mat_b.stride(0) == 1 and mat_a.stride(1) == 1
    assert scales_a.shape == (M, K // 128) and scales_b.shape == (K // 128, N // 128)
    assert bias.numel() == N and bias.is_contiguous()
    dt = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}
    out = torch.empty(M, N, dtype=out_dtype, device=mat_a.device)
    with torch.cuda.device(mat_a.device):
        stream = torch.cuda.current_stream(mat_a.device).cuda_stream
        capi.check(capi.lib.b2_gemm_e4m3_block128(
            mat_a.data_ptr(), mat_a.stride(0) if M > 1 else K, mat_b.data_ptr(), mat_b.stride(1) if N > 1 else K,
            scales_a.data_ptr(), scales_a.stride(0), scales_a.stride(1), scales_b.data_ptr(), scales_b.stride(1),
            scales_b.stride(0), bias.data_ptr(), dt[bias.dtype], out.data_ptr(), N, dt[out_dtype], M, N, K, stream))
    return out


def fp8_linear_fused(x: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor, bias=None) -> torch.Tensor:
    """W8A8-fp8 linear with dynamic per-token activation scales in ONE call of the C ABI (b2_linear_w8a8_fp8): x [M,K]
    bf16/f16, weight [K,N] e4m3 with stride(0) == 1 (the linear method's layout), weight_scale [N(,1)].  With
    capi.fused_decode() and M <= 32 the library runs it as one kernel; otherwise quantizer + GEMM through a workspace."""
    from . import capi

    M, K = x.shape
    N = weight.shape[1]
    assert weight.shape[0] == K and weight.stride(0) == 1 and x.stride(1) == 1 and weight_scale.numel() == N
    out = torch.empty(M, N, dtype=x.dtype, device=x.device)
    dt = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}
    ws_bytes = int(capi.lib.b2_linear_w8a8_fp8_workspace(M, K))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device) if ws_bytes else None
    with torch.cuda.device(x.device):
        capi.check(capi.lib.b2_linear_w8a8_fp8(x.data_ptr(), dt[x.dtype], x.stride(0) if M > 1 else K, weight.data_ptr(),
                                               weight.stride(1) if N > 1 else K, weight_scale.data_ptr(),
                                               None if bias is None else bias.data_ptr(),
                                               dt[bias.dtype] if bias is not None else capi.DTYPE_F32, out.data_ptr(), N,
                                               M, N, K, None if ws is None else ws.data_ptr(), ws_bytes,
                                               torch.cuda.current_stream(x.device).cuda_stream))
    return out


def bmm_fp8(A, B, A_scale, B_scale, dtype, out=None):
    """sgl-kernel/python/sgl_kernel/gemm.py:65-81: A [b,m,k] e4m3, B [b,k,n] e4m3 column-major, per-tensor scales."""
    if out is None:
        out = torch.empty((A.shape[0], A.shape[1], B.shape[2]), device=A.device, dtype=dtype)
    ws = _cache_buf("bmm_fp8_workspace", 1 << 20, A.device)
    _ops.bmm_fp8.default(A, B, out, A_scale, B_scale, ws, 0)
    return out


_BUFS = {}


def _cache_buf(name, nbytes, device):
    key = (name, torch.device(device))
    buf = _BUFS.get(key)
    if buf is None:
        buf = _BUFS[key] = torch.empty(nbytes, dtype=torch.uint8, device=device)
    return buf


def int8_scaled_mm_with_quant(mat1, mat2, scales2, bias, out_dtype, is_vnni=False):
    """CPU-path op name (gemm_int8.cpp:481-547) on CUDA tensors: bf16/f16 in, fused
    per-token int8 quantization + int8 GEMM + scale/bias epilogue.  mat2 is [N,K] int8."""
    return _ops.int8_scaled_mm_with_quant.default(mat1, mat2, scales2, bias, out_dtype, is_vnni)


# -------------------------------------------------------------- quantizers ----

def per_token_quant_int8_cpu(A: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """CPU-path op name (gemm_int8.cpp:382-410) on CUDA: (int8 codes [M,K], fp32 scales [M]);
    codes are signed (= the CPU op's uint8 - 128)."""
    return _ops.per_token_quant_int8_cpu.default(A)


def per_token_quant_int8(x: torch.Tensor, scale_dtype=torch.float32, cal_sum: bool = False):
    """Signature of int8_kernel.py:59-89 (the Triton quantizer the CUDA linear method calls):
    returns (x_q int8 like x, scales [..., 1]).  Arithmetic is the CPU path's (eps 1e-7,
    round-half-even), which BASELINE.json makes the bit-exact target."""
    assert x.is_contiguous()
    K = x.shape[-1]
    if cal_sum:  # int8_kernel.py:46-56 (QoQ): also the row sums of x, in x's dtype
        from . import capi

        x2 = x.view(-1, K)
        M = x2.shape[0]
        q = torch.empty_like(x, dtype=torch.int8)
        s = torch.empty(x.shape[:-1] + (1,), dtype=torch.float32, device=x.device)
        x_sum = torch.empty(x.shape[:-1], dtype=x.dtype, device=x.device)
        dt = capi.DTYPE_BF16 if x.dtype == torch.bfloat16 else capi.DTYPE_F16
        with torch.cuda.device(x.device):
            capi.check(capi.lib.b2_quant_int8_rows_sum(x2.data_ptr(), dt, K, q.data_ptr(), K, s.data_ptr(),
                                                       x_sum.data_ptr(), M, K,
                                                       torch.cuda.current_stream(x.device).cuda_stream))
        return q, (s if scale_dtype == torch.float32 else s.to(scale_dtype)), x_sum
    q, s = _ops.per_token_quant_int8_cpu.default(x.view(-1, K))
    s = s.view(x.shape[:-1] + (1,))
    if scale_dtype != torch.float32:
        s = s.to(scale_dtype)
    return q.view(x.shape), s


def sgl_per_token_quant_fp8(input: torch.Tensor, output_q: torch.Tensor, output_s: torch.Tensor) -> None:
    _ops.sgl_per_token_quant_fp8.default(input, output_q, output_s)


def sglang_per_token_quant_fp8(x: torch.Tensor, dtype: torch.dtype = fp8_dtype):
    """fp8_kernel.py:587-603."""
    assert x.is_contiguous(), "`x` is not contiguous"
    x_q = torch.empty_like(x, device=x.device, dtype=dtype)
    x_s = torch.empty(x.shape[0], 1, device=x.device, dtype=torch.float32)
    sgl_per_token_quant_fp8(x, x_q, x_s)
    return x_q, x_s


def sgl_per_token_group_quant_8bit(input, output_q, output_s, group_size, eps, fp8_min, fp8_max,
                                   scale_ue8m0=False, fuse_silu_and_mul=False, masked_m=None,
                                   enable_v2: Optional[bool] = None) -> None:
    """sgl-kernel/python/sgl_kernel/gemm.py:99-135."""
    if enable_v2 is None:
        enable_v2 = fuse_silu_and_mul or masked_m is not None
    if enable_v2:
        return _ops.sgl_per_token_group_quant_8bit_v2.default(input, output_q, output_s, group_size, eps, fp8_min,
                                                              fp8_max, scale_ue8m0, fuse_silu_and_mul, masked_m)
    assert not fuse_silu_and_mul, "only v2 support fuse_silu_and_mul"
    assert masked_m is None, "only v2 support masked_m"
    _ops.sgl_per_token_group_quant_8bit.default(input, output_q, output_s, group_size, eps, fp8_min, fp8_max,
                                                scale_ue8m0)


def sgl_per_tensor_quant_fp8(input, output_q, output_s, is_static: bool) -> None:
    """sgl-kernel/python/sgl_kernel/gemm.py:143-151."""
    _ops.sgl_per_tensor_quant_fp8.default(input, output_q, output_s, is_static)


def scaled_fp8_quant(input: torch.Tensor, scale: Optional[torch.Tensor] = None,
                     num_token_padding: Optional[int] = None, use_per_token_if_dynamic: bool = False):
    """fp8_kernel.py:1650-1684 (the sgl_kernel branch)."""
    assert input.ndim == 2, f"Expected 2D input tensor, got {input.ndim}D"
    shape = input.shape
    if num_token_padding:
        shape = (max(num_token_padding, input.shape[0]), shape[1])
    output = torch.empty(shape, device=input.device, dtype=fp8_dtype)
    if scale is None:
        if use_per_token_if_dynamic:
            scale = torch.empty((shape[0], 1), device=input.device, dtype=torch.float32)
            sgl_per_token_quant_fp8(input, output[: input.shape[0]], scale[: input.shape[0]])
        else:
            scale = torch.zeros(1, device=input.device, dtype=torch.float32)
            sgl_per_tensor_quant_fp8(input, output, scale, is_static=False)
    else:
        assert scale.numel() == 1, f"Expected scalar scale, got numel={scale.numel()}"
        sgl_per_tensor_quant_fp8(input, output, scale, is_static=True)
    return output, scale


def static_quant_fp8(x: torch.Tensor, x_s: torch.Tensor, repeat_scale: bool = False):
    """fp8_kernel.py:649-700: quantize with the given per-tensor scale; repeat_scale also returns the scale once
    per token row ([M,1]) for the per-token GEMM epilogue.  One launch of b2_quant_fp8_tensor."""
    from . import capi

    assert x.is_contiguous(), "`x` is not contiguous"
    assert x_s.numel() == 1, "only supports per-tensor scale"
    assert x_s.dtype == torch.float32 and x_s.device == x.device
    x_q = torch.empty_like(x, dtype=fp8_dtype)
    M = x.numel() // x.shape[-1]
    x_s_repeat = torch.empty((M, 1), device=x.device, dtype=torch.float32) if repeat_scale else None
    if x.numel() > 0:
        dt = capi.DTYPE_BF16 if x.dtype == torch.bfloat16 else capi.DTYPE_F16
        with torch.cuda.device(x.device):
            capi.check(capi.lib.b2_quant_fp8_tensor(x.data_ptr(), dt, x_q.data_ptr(), x_s.data_ptr(),
                                                    x_s_repeat.data_ptr() if repeat_scale else None, x.numel(), M, 1,
                                                    torch.cuda.current_stream(x.device).cuda_stream))
    return x_q, (x_s_repeat if repeat_scale else x_s)


_QMODES = {"int8": 0, "fp8": 1, "fp8_group128": 2, "fp8_group128_v2": 3}


def silu_and_mul_quant(x: torch.Tensor, mode: str = "int8", return_act: bool = False):
    """silu_and_mul (sgl-kernel/csrc/elementwise/activation.cu) fused into the activation quantizer of the following
    down_proj: x [..., 2K] (gate | up) -> (q [..., K], scales) with mode "int8" | "fp8" (per-token, scales [..., 1]) or
    "fp8_group128_v2" (1x128, M-major scales as the block-wise GEMM wants).  One kernel, one read of x."""
    from . import capi

    assert x.is_contiguous() and x.shape[-1] % 2 == 0
    K = x.shape[-1] // 2
    x2 = x.view(-1, 2 * K)
    M = x2.shape[0]
    qm = _QMODES[mode]
    assert qm in (0, 1, 3)
    q = torch.empty(x.shape[:-1] + (K,), dtype=torch.int8 if qm == 0 else fp8_dtype, device=x.device)
    if qm == 3:
        s = create_per_token_group_quant_fp8_output_scale((M, K), x.device, 128, True)
        srow, sgrp = s.stride(0), s.stride(1)
    else:
        s = torch.empty(x.shape[:-1] + (1,), dtype=torch.float32, device=x.device)
        srow, sgrp = 1, 0
    y = torch.empty(x.shape[:-1] + (K,), dtype=x.dtype, device=x.device) if return_act else None
    if M > 0:
        dt = capi.DTYPE_BF16 if x.dtype == torch.bfloat16 else capi.DTYPE_F16
        with torch.cuda.device(x.device):
            capi.check(capi.lib.b2_silu_mul_quant(x2.data_ptr(), dt, 2 * K, qm, q.data_ptr(), K, s.data_ptr(), srow, sgrp,
                                                  0, y.data_ptr() if return_act else None, K, None, 0, M, K,
                                                  torch.cuda.current_stream(x.device).cuda_stream))
    return (q, s, y) if return_act else (q, s)


def rmsnorm_quant(x: torch.Tensor, weight: torch.Tensor, eps: float = 1e-6, residual: Optional[torch.Tensor] = None,
                  mode: str = "int8", return_normed: bool = False):
    """rmsnorm / fused_add_rmsnorm (sgl-kernel/python/sgl_kernel/elementwise.py) fused into the activation quantizer
    of the following qkv_proj / gate_up_proj.  With `residual` the sum x + residual is written back to it in place
    (the reference op's contract).  mode "int8" | "fp8" | "fp8_group128".  Returns (q, scales[, normed])."""
    from . import capi

    assert x.is_contiguous() and weight.is_contiguous() and weight.dtype == x.dtype and weight.numel() == x.shape[-1]
    K = x.shape[-1]
    x2 = x.view(-1, K)
    M = x2.shape[0]
    qm = _QMODES[mode]
    assert qm in (0, 1, 2)
    if residual is not None:
        assert residual.is_contiguous() and residual.shape == x.shape and residual.dtype == x.dtype
    q = torch.empty_like(x, dtype=torch.int8 if qm == 0 else fp8_dtype)
    if qm == 2:
        s = create_per_token_group_quant_fp8_output_scale((M, K), x.device, 128, True)
        srow, sgrp = s.stride(0), s.stride(1)
    else:
        s = torch.empty(x.shape[:-1] + (1,), dtype=torch.float32, device=x.device)
        srow, sgrp = 1, 0
    y = torch.empty_like(x) if return_normed else None
    if M > 0:
        dt = capi.DTYPE_BF16 if x.dtype == torch.bfloat16 else capi.DTYPE_F16
        with torch.cuda.device(x.device):
            capi.check(capi.lib.b2_rmsnorm_quant(x2.data_ptr(), dt, K, residual.data_ptr() if residual is not None else None,
                                                 K, weight.data_ptr(), float(eps), qm, q.data_ptr(), K, s.data_ptr(), srow,
                                                 sgrp, y.data_ptr() if return_normed else None, K, M, K,
                                                 torch.cuda.current_stream(x.device).cuda_stream))
    return (q, s, y) if return_normed else (q, s)


def create_per_token_group_quant_fp8_output_scale(x_shape, device, group_size, column_major_scales: bool,
                                                  scale_tma_aligned: bool = False, scale_ue8m0: bool = False):
    """fp8_kernel.py:435-476 (fp32 scales only)."""
    if scale_ue8m0:
        raise NotImplementedError("UE8M0 group scales are out of scope (SURVEY 8a row a8)")
    if column_major_scales:
        if scale_tma_aligned:
            aligned = (x_shape[-2] + 3) // 4 * 4
            return torch.empty(x_shape[:-2] + (x_shape[-1] // group_size, aligned), device=device,
                               dtype=torch.float32).transpose(-1, -2)[: x_shape[-2], :]
        return torch.empty((x_shape[-1] // group_size,) + tuple(x_shape[:-1]), device=device,
                           dtype=torch.float32).permute(-1, -2)
    return torch.empty(tuple(x_shape[:-1]) + (x_shape[-1] // group_size,), device=device, dtype=torch.float32)


def sglang_per_token_group_quant_fp8(x: torch.Tensor, group_size: int, eps: float = 1e-10,
                                     column_major_scales: bool = False, scale_tma_aligned: bool = False,
                                     scale_ue8m0: bool = False, fuse_silu_and_mul: bool = False,
                                     masked_m: Optional[torch.Tensor] = None, enable_v2: Optional[bool] = None):
    """fp8_kernel.py:479-541."""
    assert x.shape[-1] % group_size == 0, "the last dimension of `x` cannot be divisible by `group_size`"
    assert x.is_contiguous(), "`x` is not contiguous"
    out_shape = (*x.shape[:-1], x.shape[-1] // (2 if fuse_silu_and_mul else 1))
    x_q = torch.empty(out_shape, device=x.device, dtype=fp8_dtype)
    x_s = create_per_token_group_quant_fp8_output_scale(out_shape, x.device, group_size, column_major_scales,
                                                        scale_tma_aligned, scale_ue8m0)
    if x.shape[0] > 0:
        sgl_per_token_group_quant_8bit(x, x_q, x_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0,
                                       fuse_silu_and_mul, masked_m, enable_v2=enable_v2)
    return x_q, x_s


per_token_group_quant_fp8 = sglang_per_token_group_quant_fp8


# ------------------------------------------------------- fake (meta) kernels ----

def _register_fakes() -> None:
    from torch.library import register_fake

    def _mm_fake(mat_a, mat_b, scales_a, scales_b, out_dtype, bias=None):
        return mat_a.new_empty((mat_a.shape[0], mat_b.shape[1]), dtype=out_dtype)

    def _bw_fake(mat_a, mat_b, scales_a, scales_b, out_dtype):
        return mat_a.new_empty((mat_a.shape[0], mat_b.shape[1]), dtype=out_dtype)

    def _quant_fake(A):
        return A.new_empty(A.shape, dtype=torch.int8), A.new_empty((A.shape[0],), dtype=torch.float32)

    def _mm_cpu_fake(mat1, mat2, scales1, scales2, bias, out_dtype, is_vnni):
        return mat1.new_empty((mat1.shape[0], mat2.shape[0]), dtype=out_dtype)

    def _wq_fake(mat1, mat2, scales2, bias, out_dtype, is_vnni):
        return mat1.new_empty((mat1.shape[0], mat2.shape[0]), dtype=out_dtype)

    def _void3(input, output_q, output_s):
        return None

    def _void8(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0):
        return None

    def _void10(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0, fuse_silu_and_mul, masked_m):
        return None

    def _void4(input, output_q, output_s, is_static):
        return None

    def _void7(A, B, D, A_scale, B_scale, workspace_buffer, cublas_handle):
        return None

    for name, fn in (
        ("sgl_kernel::sgl_per_token_group_quant_8bit_v2", _void10), ("sgl_kernel::sgl_per_tensor_quant_fp8", _void4),
        ("sgl_kernel::bmm_fp8", _void7),
        ("sgl_kernel::int8_scaled_mm", _mm_fake), ("sgl_kernel::fp8_scaled_mm", _mm_fake),
        ("sgl_kernel::fp8_blockwise_scaled_mm", _bw_fake), ("sgl_kernel::per_token_quant_int8_cpu", _quant_fake),
        ("sgl_kernel::int8_scaled_mm_cpu", _mm_cpu_fake), ("sgl_kernel::int8_scaled_mm_with_quant", _wq_fake),
        ("sgl_kernel::sgl_per_token_quant_fp8", _void3), ("sgl_kernel::sgl_per_token_group_quant_8bit", _void8),
    ):
        try:
            register_fake(name)(fn)
        except RuntimeError:
            pass  # a caller (sglang itself) registered one already


_register_fakes()
