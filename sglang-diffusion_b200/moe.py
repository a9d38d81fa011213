"""W8A8-int8 mixture of experts on the grouped tcgen05 GEMM.

Mirrors, with the same names / argument meaning / weight layouts:
  W8A8Int8MoEMethod   python/sglang/srt/layers/quantization/w8a8_int8.py:234-378
                      (w13_weight [E, 2I, H] int8, w2_weight [E, H, I] int8, per-channel fp32 scales [E, 2I, 1], [E, H, 1])
  fused_experts       the int8 W8A8 branch of torch.ops.sgl_kernel.fused_experts_cpu
                      (sgl-kernel/csrc/cpu/moe.cpp fused_experts_cpu -> moe_int8.cpp:585-847): quantize, gate/up GEMM
                      over expert-sorted row blocks, silu_and_mul, quantize, down GEMM, weighted sum over top-k.

Seven launches of this library per call, no torch arithmetic: block alignment, gather + quantize, grouped GEMM,
fused silu_and_mul + quantize, grouped GEMM, combine.  (shared_expert_cpu, moe_int8.cpp:849, is a plain two-layer
MLP: the dense ops already cover it.)
"""
from __future__ import annotations

from typing import Optional

import torch

from . import capi, ops


def _dt(t: torch.Tensor) -> int:
    return capi.DTYPE_BF16 if t.dtype == torch.bfloat16 else capi.DTYPE_F16


def moe_align_block128(topk_ids: torch.Tensor, num_experts: int):
    """(sorted_ids [P], tile_expert [P/128], inv_pos [n], num_post_pad [1]) -- b2_moe_align_block128."""
    assert topk_ids.dtype == torch.int32 and topk_ids.is_contiguous()
    n = topk_ids.numel()
    P = int(capi.lib.b2_moe_max_padded(n, num_experts))
    dev = topk_ids.device
    sorted_ids = torch.empty(P, dtype=torch.int32, device=dev)
    tile_expert = torch.empty(P // 128, dtype=torch.int32, device=dev)
    inv_pos = torch.empty(n, dtype=torch.int32, device=dev)
    num_post_pad = torch.empty(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        capi.check(capi.lib.b2_moe_align_block128(topk_ids.data_ptr(), n, num_experts, sorted_ids.data_ptr(),
                                                  tile_expert.data_ptr(), inv_pos.data_ptr(), num_post_pad.data_ptr(), P,
                                                  torch.cuda.current_stream(dev).cuda_stream))
    return sorted_ids, tile_expert, inv_pos, num_post_pad


def grouped_int8_mm(a_q: torch.Tensor, a_s: torch.Tensor, w: torch.Tensor, w_s: torch.Tensor,
                    tile_expert: torch.Tensor, out_dtype: torch.dtype) -> torch.Tensor:
    """a_q [P, K] int8 expert-sorted rows, w [E, N, K] int8, w_s [E, N(,1)] fp32 -> [P, N]."""
    P, K = a_q.shape
    E, N, K2 = w.shape
    assert K2 == K and w.is_contiguous() and a_q.is_contiguous() and w_s.numel() == E * N and w_s.is_contiguous()
    out = torch.empty(P, N, dtype=out_dtype, device=a_q.device)
    with torch.cuda.device(a_q.device):
        capi.check(capi.lib.b2_gemm_s8_grouped(a_q.data_ptr(), K, w.data_ptr(), K, a_s.data_ptr(), w_s.data_ptr(),
                                               tile_expert.data_ptr(), E, out.data_ptr(), N,
                                               capi.DTYPE_BF16 if out_dtype == torch.bfloat16 else capi.DTYPE_F16,
                                               P, N, K, torch.cuda.current_stream(a_q.device).cuda_stream))
    return out


def fused_experts(hidden_states: torch.Tensor, w1: torch.Tensor, w2: torch.Tensor, topk_weights: torch.Tensor,
                  topk_ids: torch.Tensor, w1_scale: torch.Tensor, w2_scale: torch.Tensor,
                  inplace: bool = False) -> torch.Tensor:
    """hidden_states [T, H] bf16/f16; w1 [E, 2I, H] int8 (gate rows, then up rows); w2 [E, H, I] int8;
    topk_weights [T, topk] fp32; topk_ids [T, topk] int32.  Returns [T, H]."""
    T, H = hidden_states.shape
    E, I2, _ = w1.shape
    I = I2 // 2
    topk = topk_ids.shape[1]
    dev = hidden_states.device
    x = hidden_states.contiguous()
    ids = topk_ids.to(torch.int32).contiguous()
    tw = topk_weights.to(torch.float32).contiguous()
    sorted_ids, tile_expert, inv_pos, _ = moe_align_block128(ids, E)
    P = sorted_ids.numel()
    with capi.static_weights(), torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        a_q = torch.empty(P, H, dtype=torch.int8, device=dev)
        a_s = torch.empty(P, dtype=torch.float32, device=dev)
        capi.check(capi.lib.b2_quant_int8_rows_gather(x.data_ptr(), _dt(x), H, sorted_ids.data_ptr(), topk, T * topk,
                                                      a_q.data_ptr(), H, a_s.data_ptr(), P, H, stream))
        ic1 = grouped_int8_mm(a_q, a_s, w1, w1_scale, tile_expert, x.dtype)          # [P, 2I]
        h_q, h_s = ops.silu_and_mul_quant(ic1, "int8")                               # [P, I], [P, 1]
        ic2 = grouped_int8_mm(h_q, h_s.view(-1), w2, w2_scale, tile_expert, x.dtype)  # [P, H]
        out = hidden_states if inplace else torch.empty(T, H, dtype=x.dtype, device=dev)
        capi.check(capi.lib.b2_moe_combine(ic2.data_ptr(), H, inv_pos.data_ptr(), tw.data_ptr(), out.data_ptr(), H, T,
                                           topk, H, _dt(x), stream))
    return out


class W8A8Int8MoEMethod:
    """w8a8_int8.py:234-378: per-channel int8 expert weights, dynamic per-token int8 activations."""

    def create_weights(self, layer: torch.nn.Module, num_experts: int, hidden_size: int,
                       intermediate_size_per_partition: int, params_dtype: torch.dtype, device=None, **_):
        layer.w13_weight = torch.nn.Parameter(
            torch.empty(num_experts, 2 * intermediate_size_per_partition, hidden_size, dtype=torch.int8, device=device),
            requires_grad=False)
        layer.w2_weight = torch.nn.Parameter(
            torch.empty(num_experts, hidden_size, intermediate_size_per_partition, dtype=torch.int8, device=device),
            requires_grad=False)
        layer.w13_weight_scale = torch.nn.Parameter(
            torch.ones(num_experts, 2 * intermediate_size_per_partition, 1, dtype=torch.float32, device=device),
            requires_grad=False)
        layer.w2_weight_scale = torch.nn.Parameter(
            torch.ones(num_experts, hidden_size, 1, dtype=torch.float32, device=device), requires_grad=False)
        layer.w13_input_scale = None
        layer.w2_input_scale = None

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        pass  # [E, N, K] K-contiguous is what the grouped GEMM's tensor map wants: no repacking (w8a8_int8.py:319-326)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor,
              apply_router_weight_on_input: bool = False) -> torch.Tensor:
        assert not apply_router_weight_on_input, "apply_router_weight_on_input is not built"
        return fused_experts(x, layer.w13_weight, layer.w2_weight, topk_weights, topk_ids, layer.w13_weight_scale,
                             layer.w2_weight_scale)
