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
    stream = torch.cuda.current_stream(mat_a.device).cuda_stream
    capi.check(fn(mat_a.data_ptr(), mat_a.stride(0), mat_b.data_ptr(), mat_b.stride(1), scales_a.data_ptr(),
                  scales_b.data_ptr(), None if bias is None else bias.data_ptr(),
                  dt[bias.dtype] if bias is not None else capi.DTYPE_F32, out.data_ptr(), out.stride(0),
                  dt[out.dtype], M, N, K, stream))
    return out


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
    if cal_sum:
        raise NotImplementedError("per_token_quant_int8(cal_sum=True) (QoQ row sums) is out of scope")
    assert x.is_contiguous()
    K = x.shape[-1]
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
                                   scale_ue8m0=False) -> None:
    _ops.sgl_per_token_group_quant_8bit.default(input, output_q, output_s, group_size, eps, fp8_min, fp8_max,
                                                scale_ue8m0)


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
    if fuse_silu_and_mul or masked_m is not None:
        raise NotImplementedError("fused silu_and_mul / masked_m group quant is a 'next' row (SURVEY 8f rank 1)")
    x_q = torch.empty(x.shape, device=x.device, dtype=fp8_dtype)
    x_s = create_per_token_group_quant_fp8_output_scale(x.shape, x.device, group_size, column_major_scales,
                                                        scale_tma_aligned, scale_ue8m0)
    if x.shape[0] > 0:
        sgl_per_token_group_quant_8bit(x, x_q, x_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0)
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

    for name, fn in (
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
