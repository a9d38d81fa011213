"""TEST INFRASTRUCTURE ONLY -- ctypes/torch front-end of oracle/restate.c.

Never imported by the product package (sglang-diffusion_b200/).  Used by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.

The functions mirror the reference's CPU operator names and argument meaning
(sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:415-447) with two additions the
reference never exposes: the int32 accumulators, and the W8A8 fp8 restatements of
the reference's CUDA ops.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "restate.c")
OUT_DIR = os.path.join(HERE, "_build")
OUT_SO = os.path.join(OUT_DIR, "liboracle_restate.so")
_LIB = None


def build(force: bool = False) -> str:
    """gcc -O3 -fopenmp -ffp-contract=off, generic x86-64-v3 so the .so runs on any host."""
    os.makedirs(OUT_DIR, exist_ok=True)
    if not force and os.path.isfile(OUT_SO) and os.path.getmtime(OUT_SO) >= os.path.getmtime(SRC):
        return OUT_SO
    cmd = [
        "/usr/bin/gcc", "-O3", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
        "-march=x86-64-v3", "-D_GNU_SOURCE", SRC, "-o", OUT_SO, "-lm",
    ]
    subprocess.check_call(cmd)
    return OUT_SO


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.isfile(OUT_SO):
            build()
        _LIB = ctypes.CDLL(OUT_SO)
        _LIB.orc_e4m3_to_f32.restype = ctypes.c_float
        _LIB.orc_e4m3_to_f32.argtypes = [ctypes.c_uint8]
        _LIB.orc_f32_to_e4m3.restype = ctypes.c_uint8
        _LIB.orc_f32_to_e4m3.argtypes = [ctypes.c_float]
    return _LIB


def _p(t: torch.Tensor | None):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _i(v: int):
    return ctypes.c_int64(int(v))


def _chk_cpu(*ts):
    for t in ts:
        if t is not None:
            assert t.device.type == "cpu", "oracle runs on CPU tensors only"


# ------------------------------------------------------------------ int8 ----

def per_token_quant_int8_cpu(A: torch.Tensor):
    """(u8 codes = q+128 [M,K], scales f32 [M]) -- gemm_int8.cpp:382-410, vec.h:190-228."""
    _chk_cpu(A)
    assert A.dtype == torch.bfloat16 and A.dim() == 2 and A.stride(1) == 1
    M, K = A.shape
    q = torch.empty(M, K, dtype=torch.uint8)
    s = torch.empty(M, dtype=torch.float32)
    lib().orc_quant_int8_rows(_p(A), _i(A.stride(0)), _p(q), _p(s), _i(M), _i(K))
    return q, s


def int8_acc(Aq_u8: torch.Tensor, W: torch.Tensor) -> torch.Tensor:
    """int32 accumulators of the reference's u8*s8 - compensation kernel (exact s8.s8 dot)."""
    _chk_cpu(Aq_u8, W)
    assert Aq_u8.dtype == torch.uint8 and W.dtype == torch.int8
    Aq_u8, W = Aq_u8.contiguous(), W.contiguous()
    M, K = Aq_u8.shape
    N = W.shape[0]
    assert W.shape[1] == K
    acc = torch.empty(M, N, dtype=torch.int32)
    lib().orc_gemm_u8s8_acc(_p(Aq_u8), _p(W), _p(acc), _i(M), _i(N), _i(K))
    return acc


def int8_epilogue(acc, As, Bs, bias):
    _chk_cpu(acc, As, Bs, bias)
    M, N = acc.shape
    As = As.reshape(-1).float().contiguous()
    Bs = Bs.reshape(-1).float().contiguous()
    b = None if bias is None else bias.reshape(-1).float().contiguous()
    assert As.numel() == M and Bs.numel() == N
    out = torch.empty(M, N, dtype=torch.bfloat16)
    lib().orc_int8_epilogue(_p(acc.contiguous()), _p(As), _p(Bs), _p(b), _p(out), _i(M), _i(N))
    return out


def int8_scaled_mm_cpu(mat1_u8, mat2, scales1, scales2, bias, out_dtype=torch.bfloat16, is_vnni=False):
    """gemm_int8.cpp:422-478.  mat2 is the UNPACKED [N,K] int8 weight (is_vnni ignored)."""
    assert out_dtype == torch.bfloat16, "reference CPU path is bf16-only in practice (SURVEY fact 7)"
    return int8_epilogue(int8_acc(mat1_u8, mat2), scales1, scales2, bias)


def int8_scaled_mm_with_quant(mat1, mat2, scales2, bias, out_dtype=torch.bfloat16, is_vnni=False,
                              return_parts=False):
    """gemm_int8.cpp:481-547 (quantize rows, then the int8 GEMM + epilogue)."""
    assert out_dtype == torch.bfloat16
    q, s = per_token_quant_int8_cpu(mat1)
    acc = int8_acc(q, mat2)
    out = int8_epilogue(acc, s, scales2, bias)
    if return_parts:
        return out, q, s, acc
    return out


# ------------------------------------------------------------------- fp8 ----

def e4m3_decode_table() -> torch.Tensor:
    L = lib()
    return torch.tensor([L.orc_e4m3_to_f32(i) for i in range(256)], dtype=torch.float32)


def per_token_quant_fp8(x: torch.Tensor):
    """(e4m3 codes as uint8 [M,K], scales f32 [M]) -- per_token_quant_fp8.cu:17-121, IEEE arithmetic."""
    _chk_cpu(x)
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    q = torch.empty(M, K, dtype=torch.uint8)
    s = torch.empty(M, dtype=torch.float32)
    lib().orc_quant_fp8_rows(_p(x), _i(x.stride(0)), _p(q), _p(s), _i(M), _i(K))
    return q, s


def per_token_group_quant_fp8(x: torch.Tensor, group_size=128, eps=1e-10, column_major_scales=True):
    """(codes uint8 [M,K], scales f32 logical [M,K/G]; M-major storage when column_major_scales)
    -- per_token_group_quant_8bit.cu:19-114."""
    _chk_cpu(x)
    assert x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    assert K % group_size == 0
    ng = K // group_size
    q = torch.empty(M, K, dtype=torch.uint8)
    if column_major_scales:
        store = torch.empty(ng, M, dtype=torch.float32)
        stride = M
    else:
        store = torch.empty(M, ng, dtype=torch.float32)
        stride = 0
    lib().orc_quant_fp8_group(_p(x), _i(x.stride(0)), _p(q), _p(store), _i(stride), _i(M), _i(K),
                              _i(group_size), ctypes.c_float(eps), ctypes.c_float(-448.0), ctypes.c_float(448.0))
    s = store.t() if column_major_scales else store
    return q, s


def fp8_scaled_mm(Aq_u8, W_u8, As, Bs, bias=None) -> torch.Tensor:
    """fp32 result of As*(Bs*(A.W^T)) [+bias] with exact products, double accumulation."""
    _chk_cpu(Aq_u8, W_u8, As, Bs, bias)
    Aq_u8, W_u8 = Aq_u8.contiguous(), W_u8.contiguous()
    M, K = Aq_u8.shape
    N = W_u8.shape[0]
    As = As.reshape(-1).float().contiguous()
    Bs = Bs.reshape(-1).float().contiguous()
    b = None if bias is None else bias.reshape(-1).float().contiguous()
    out = torch.empty(M, N, dtype=torch.float32)
    lib().orc_gemm_e4m3(_p(Aq_u8), _p(W_u8), _p(As), _p(Bs), _p(b), _p(out), _i(M), _i(N), _i(K))
    return out


def fp8_blockwise_scaled_mm(Aq_u8, W_u8, sfa, sfb) -> torch.Tensor:
    """sfa logical [M,K/128], sfb logical [N/128,K/128] (any strides; copied to the C layout)."""
    _chk_cpu(Aq_u8, W_u8, sfa, sfb)
    Aq_u8, W_u8 = Aq_u8.contiguous(), W_u8.contiguous()
    M, K = Aq_u8.shape
    N = W_u8.shape[0]
    sfa_c = sfa.float().t().contiguous()  # [K/128, M]  -> element (m,kb) at kb*M + m
    sfb_c = sfb.float().contiguous()      # [N/128, K/128] -> element (nb,kb) at nb*ld + kb
    out = torch.empty(M, N, dtype=torch.float32)
    lib().orc_gemm_e4m3_block128(_p(Aq_u8), _p(W_u8), _p(sfa_c), _i(M), _p(sfb_c), _i(sfb_c.shape[1]), _p(out),
                                 _i(M), _i(N), _i(K))
    return out


# ------------------------------------------- producer fusion / per-tensor (8f) ----
# Restated in torch fp32 on CPU (every step an IEEE fp32 operation, same order as the CUDA sources named).

def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """y = T(silu(f32(gate))) * up in T -- per_token_group_quant_8bit_v2.cu:330-341.  The sm_100 source computes silu
    as 0.5x(1 + tanh.approx(0.5x)) (:32-40), a hardware approximation this exact-tanh restatement can only bound:
    tolerance-pinned (<= 1 ulp of T on the activation, see tests)."""
    _chk_cpu(x)
    K = x.shape[-1] // 2
    g, u = x[..., :K].float(), x[..., K:]
    half = g * 0.5
    act = (half * (1.0 + torch.tanh(half))).to(x.dtype)
    return (act.float() * u.float()).to(x.dtype)


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: torch.Tensor | None = None):
    """flashinfer norm.cuh FusedAddRMSNormKernel / RMSNormKernel as called by
    sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu: s = f32(x) + f32(res); res <- T(s);
    y = T((s * rsqrt(mean(s^2) + eps)) * w).  Returns (y, new_residual or None).  The sum of squares is reduced in a
    different order on the GPU and rsqrt is 2-ulp there: tolerance-pinned."""
    _chk_cpu(x, weight, residual)
    s = x.float() if residual is None else x.float() + residual.float()
    new_res = None if residual is None else s.to(x.dtype)
    rcp = torch.rsqrt((s.double() * s.double()).sum(-1, keepdim=True).float() / x.shape[-1] + eps)
    y = ((s * rcp) * weight.float()).to(x.dtype)
    return y, new_res


def per_token_group_quant_fp8_v2(y: torch.Tensor, group_size=128):
    """(codes uint8 [M,K], scales f32 [M,K/G]) -- per_token_group_quant_8bit_v2.cu:69-81,343-372, fp32 scales:
    amax = max(1e-10, max|y|); s = amax * f32(1/448); q = e4m3_rn(clamp(y * (448/amax), +-448))."""
    _chk_cpu(y)
    M, K = y.shape
    g = y.float().view(M, K // group_size, group_size)
    amax = g.abs().amax(-1, keepdim=True).clamp_min(1e-10)
    s = amax * torch.tensor(1.0 / 448.0, dtype=torch.float32)
    ys = torch.tensor(448.0, dtype=torch.float32) / amax
    q = (g * ys).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8).view(M, K)
    return q, s.squeeze(-1)


def per_tensor_quant_fp8(x: torch.Tensor, scale: torch.Tensor | None = None):
    """(codes uint8 like x, scale f32 [1]) -- per_tensor_quant_fp8.cu:11-96: dynamic s = max|x|/448;
    q = e4m3_rn(clamp(x * (1/s), +-448)).  (The reference build's --use_fast_math turns 1/s into rcp.approx; the IEEE
    quotient is the documented arithmetic and what the library computes.)"""
    _chk_cpu(x, scale)
    if scale is None:
        scale = (x.float().abs().max() / 448.0).reshape(1)
    inv = torch.tensor(1.0, dtype=torch.float32) / scale.float().reshape(1)
    q = (x.float() * inv).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8)
    return q, scale.float().reshape(1)


def fused_experts_int8(x, w13, w2, topk_weights, topk_ids, w13_scale, w2_scale):
    """fused_experts_int8_kernel_impl (moe_int8.cpp:585-847) restated per (token, k): quantize x rows (bit-exact
    restatement), h = int8 linear with the expert's gate/up weight (bit-exact), act = silu_and_mul (exact tanh:
    tolerance-pinned, see silu_and_mul), quantize, int8 linear with the expert's down weight, fp32 weighted sum over k
    in ascending order, one rounding to bf16."""
    _chk_cpu(x, w13, w2, topk_weights, topk_ids, w13_scale, w2_scale)
    T, H = x.shape
    E = w13.shape[0]
    out = torch.zeros(T, H, dtype=torch.float32)
    ids = topk_ids.long()
    for e in range(E):
        tok, kk = (ids == e).nonzero(as_tuple=True)
        if tok.numel() == 0:
            continue
        h = int8_scaled_mm_with_quant(x[tok].contiguous(), w13[e].contiguous(), w13_scale[e].reshape(-1), None)
        act = silu_and_mul(h)
        y = int8_scaled_mm_with_quant(act.contiguous(), w2[e].contiguous(), w2_scale[e].reshape(-1), None)
        out.index_put_((tok, ), topk_weights[tok, kk].float()[:, None] * y.float(), accumulate=True)
    return out.to(x.dtype)


if __name__ == "__main__":
    print(build(force=True))
