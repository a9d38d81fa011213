"""Host-side mirror of the reference's W8A8 linear methods and their tensor-parallel callers.

Mirrors, with the same names / argument meaning / weight layouts:
  W8A8Int8LinearMethod  python/sglang/srt/layers/quantization/w8a8_int8.py:155-231
  W8A8Fp8LinearMethod   python/sglang/srt/layers/quantization/w8a8_fp8.py:103-194
  apply_fp8_linear      python/sglang/srt/layers/quantization/fp8_utils.py:1194-1276 (dynamic per-token branch)
  cutlass_w8a8_block_fp8_linear_with_fallback   fp8_utils.py:518-548
  QuantLinear           the slice of ColumnParallelLinear / RowParallelLinear.forward that touches
                        the path (python/sglang/srt/layers/linear.py:448-456, 1467-1496):
                        quant_method.apply(...) then, for a row-parallel shard, one in-place
                        all-reduce of the [M,N] output with the bias added on tp_rank 0 only.

`fused=True` on the int8 method routes through the single entry
torch.ops.sgl_kernel.int8_scaled_mm_with_quant (the name the reference's CPU backend calls,
w8a8_int8.py:207-215), which runs quantizer + GEMM back to back inside the library with
the quantized activations kept in an L2-resident workspace.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from . import capi, ops


# --------------------------------------------------------------- int8 W8A8 ----

class W8A8Int8LinearMethod:
    """Per-channel int8 weights [N,K] + fp32 scales [N,1]; dynamic per-token int8 activations."""

    def __init__(self, fused: bool = False):
        self.fused = fused

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, device=None, **_):
        n = sum(output_partition_sizes)
        layer.logical_widths = output_partition_sizes
        layer.weight = torch.nn.Parameter(
            torch.empty(n, input_size_per_partition, dtype=torch.int8, device=device), requires_grad=False)
        layer.weight_scale = torch.nn.Parameter(
            torch.empty(n, 1, dtype=torch.float32, device=device), requires_grad=False)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        # w8a8_int8.py:160-168: a transposed VIEW ([K,N], stride(0)==1) -- no repacking on GPU;
        # both operands are already K-major, which is what the tcgen05 smem descriptors want.
        layer.weight = torch.nn.Parameter(layer.weight.t(), requires_grad=False)
        layer.weight_scale = torch.nn.Parameter(layer.weight_scale.data, requires_grad=False)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None):
        # layer.weight is a loaded parameter, never the output of the kernel before the GEMM: decode-shape GEMMs
        # may fetch it under the quantizer's tail (b2_set_weight_prefetch, include/b200_w8a8.h)
        with capi.static_weights():
            if self.fused or (x.is_cuda and x.numel() // x.shape[-1] <= 32 and x.is_contiguous()):
                # decode shapes (M <= 32) always take the single entry (quantizer + GEMM chained inside the library)
                x2d = x.view(-1, x.shape[-1])
                out = ops.int8_scaled_mm_with_quant(x2d, layer.weight.t(), layer.weight_scale, bias, x.dtype, False)
                return out.view(*x.shape[:-1], layer.weight.shape[1])
            x_q, x_scale = ops.per_token_quant_int8(x)
            x_q_2d = x_q.view(-1, x_q.shape[-1])
            x_scale_2d = x_scale.view(-1, x_scale.shape[-1])
            output_shape = [*x_q.shape[:-1], layer.weight.shape[1]]
            output = ops.int8_scaled_mm(x_q_2d, layer.weight, x_scale_2d, layer.weight_scale, out_dtype=x.dtype,
                                        bias=bias)
            return output.view(output_shape)


# ---------------------------------------------------------------- fp8 W8A8 ----

def apply_fp8_linear(input: torch.Tensor, weight: torch.Tensor, weight_scale: torch.Tensor,
                     input_scale: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
                     **_ignored) -> torch.Tensor:
    """fp8_utils.py:1194-1276 (weight is [K,N] with stride(0)==1): dynamic per-token activation scales, or -- with
    `input_scale` -- the static per-tensor branch (:1232-1239), whose scale is broadcast to one value per token so
    that the same per-token x per-channel GEMM epilogue serves both."""
    input_2d = input.view(-1, input.shape[-1])
    output_shape = [*input.shape[:-1], weight.shape[1]]
    assert weight_scale.numel() == weight.shape[1], "per-channel weight scales expected"
    if input_scale is not None:
        assert input_scale.numel() == 1
        qinput, x_scale = ops.static_quant_fp8(input_2d.contiguous(), input_scale, repeat_scale=True)
    elif input_2d.is_cuda and input_2d.shape[0] <= 32 and input_2d.is_contiguous() and weight.stride(0) == 1:
        # decode shapes: the single C-ABI entry (b2_linear_w8a8_fp8: quantizer + GEMM chained inside the library)
        return ops.fp8_linear_fused(input_2d, weight, weight_scale, bias).view(*output_shape)
    else:
        qinput, x_scale = ops.sglang_per_token_quant_fp8(input_2d)
    output = ops.fp8_scaled_mm(qinput, weight, x_scale, weight_scale, out_dtype=input.dtype, bias=bias)
    return output.view(*output_shape)


class W8A8Fp8LinearMethod:
    """Per-channel e4m3 weights [N,K] + fp32 scales [N,1]; dynamic per-token e4m3 activations."""

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size,
                       params_dtype, device=None, **_):
        n = sum(output_partition_sizes)
        layer.logical_widths = output_partition_sizes
        layer.weight = torch.nn.Parameter(
            torch.empty(n, input_size_per_partition, dtype=torch.float8_e4m3fn, device=device), requires_grad=False)
        layer.weight_scale = torch.nn.Parameter(
            torch.empty(n, 1, dtype=torch.float32, device=device), requires_grad=False)

    def process_weights_after_loading(self, layer) -> None:
        layer.weight = torch.nn.Parameter(layer.weight.t(), requires_grad=False)  # w8a8_fp8.py:120
        layer.weight_scale = torch.nn.Parameter(layer.weight_scale.data, requires_grad=False)

    def apply(self, layer, x: torch.Tensor, bias: Optional[torch.Tensor] = None):
        with capi.static_weights():  # see W8A8Int8LinearMethod.apply
            return apply_fp8_linear(x, layer.weight, layer.weight_scale,
                                    input_scale=getattr(layer, "input_scale", None), bias=bias)


def cutlass_w8a8_block_fp8_linear_with_fallback(input: torch.Tensor, weight: torch.Tensor, block_size: List[int],
                                                weight_scale: torch.Tensor,
                                                input_scale: Optional[torch.Tensor] = None,
                                                bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp8_utils.py:518-548.  weight [N,K] e4m3, weight_scale [N/128, K/128]."""
    assert input_scale is None
    assert list(block_size) == [128, 128], "only 128x128 weight blocks are built"
    if weight.shape[0] % 128 != 0 or weight.shape[1] % 128 != 0:
        raise NotImplementedError("block-fp8 shapes off the 128 grid take the reference's Triton fallback (out of scope)")
    input_2d = input.view(-1, input.shape[-1])
    output_shape = [*input.shape[:-1], weight.shape[0]]
    q_input, x_scale = ops.per_token_group_quant_fp8(input_2d, block_size[1], column_major_scales=True)
    # the reference adds the bias with a second kernel (fp8_utils.py:546-547); here it rides in the GEMM epilogue
    output = ops.fp8_blockwise_scaled_mm_bias(q_input, weight.T, x_scale, weight_scale.T, input_2d.dtype, bias)
    return output.view(*output_shape)


class W8A8BlockFp8LinearMethod:
    """Block-quantized e4m3 weights [N,K] + weight_scale_inv [N/128, K/128]
    (multimodal_gen/runtime/layers/quantization/fp8.py:459-488)."""

    block_size = [128, 128]

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size,
                       params_dtype, device=None, **_):
        n = sum(output_partition_sizes)
        layer.weight = torch.nn.Parameter(
            torch.empty(n, input_size_per_partition, dtype=torch.float8_e4m3fn, device=device), requires_grad=False)
        layer.weight_scale_inv = torch.nn.Parameter(
            torch.empty((n + 127) // 128, (input_size_per_partition + 127) // 128, dtype=torch.float32,
                        device=device), requires_grad=False)

    def process_weights_after_loading(self, layer) -> None:
        pass

    def apply(self, layer, x, bias=None):
        return cutlass_w8a8_block_fp8_linear_with_fallback(x, layer.weight, self.block_size, layer.weight_scale_inv,
                                                           None, bias)


# ------------------------------------------------------ TP caller (linear.py) ----

def shard_rows(M: int, rank: int, world_size: int):
    """Contiguous row block [start, stop) of rank `rank` when M token rows are split over
    `world_size` data-parallel ranks (weights replicated, no collective: per-token scales make
    every output row depend on its own input row only -- SURVEY 8e).  The first M % world_size
    ranks take one extra row, so block sizes differ by at most one and concatenate in rank order."""
    assert 0 <= rank < world_size and M >= 0
    base, extra = divmod(M, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class QuantLinear(torch.nn.Module):
    """The part of Column/RowParallelLinear.forward that is on the path.

    parallel = "replicated" | "column" | "row".  Column: W split on N, no collective
    (gather_output is never set inside transformer blocks).  Row: W split on K, the input is
    already K-sharded, per-token scales are local to the shard, partial outputs are summed
    with ONE all-reduce; bias only on tp_rank 0 (linear.py:1480).
    """

    def __init__(self, input_size: int, output_size: int, quant_method, bias: bool = False,
                 params_dtype: torch.dtype = torch.bfloat16, parallel: str = "replicated",
                 tp_rank: int = 0, tp_size: int = 1, tp_group=None, device=None):
        super().__init__()
        assert parallel in ("replicated", "column", "row")
        self.parallel, self.tp_rank, self.tp_size, self.tp_group = parallel, tp_rank, tp_size, tp_group
        self.input_size, self.output_size = input_size, output_size
        k_local = input_size // tp_size if parallel == "row" else input_size
        n_local = output_size // tp_size if parallel == "column" else output_size
        self.quant_method = quant_method
        quant_method.create_weights(self, k_local, [n_local], input_size, output_size, params_dtype, device=device)
        self.bias = (torch.nn.Parameter(torch.zeros(n_local, dtype=params_dtype, device=device), requires_grad=False)
                     if bias else None)

    def shard_of(self, full: torch.Tensor, dim_n: int = 0, dim_k: int = 1) -> torch.Tensor:
        """Slice a full [N,K]-shaped (or [N,...]-shaped) checkpoint tensor for this rank."""
        if self.parallel == "column":
            n = full.shape[dim_n] // self.tp_size
            return full.narrow(dim_n, self.tp_rank * n, n)
        if self.parallel == "row" and full.dim() > dim_k and full.shape[dim_k] > 1:
            k = full.shape[dim_k] // self.tp_size
            return full.narrow(dim_k, self.tp_rank * k, k)
        return full

    def use_symmetric_all_reduce(self, comm) -> None:
        """Route decode-sized row-parallel reductions through `comm` (sglang_diffusion_b200.comm.SymmAllReduce):
        the GEMM writes its partial output into the symmetric buffer and b2_allreduce_symm sums it over NVLink
        peer memory; messages that do not fit keep the NCCL all-reduce.  (The reference makes the same split
        between its custom all-reduce and NCCL, custom_all_reduce.py:328-346.)"""
        self.symm_comm = comm

    def _row_parallel_symm(self, x: torch.Tensor, bias) -> Optional[torch.Tensor]:
        comm = getattr(self, "symm_comm", None)
        if comm is None or not x.is_cuda or not isinstance(self.quant_method, (W8A8Int8LinearMethod, W8A8Fp8LinearMethod)):
            return None
        x2d = x.view(-1, x.shape[-1])
        M, N = x2d.shape[0], self.weight.shape[1]
        # fused push moves world x the message into every rank (an all-gather of partials): measured ahead of the
        # separate NVSwitch two-shot reduction only while that stays small (profiles/r02j_bench_{2,8}gpu.json:
        # M = 64, N = 8192 on 8 ranks: 28.0 us pushed vs 20.9 us two-shot vs 37.4 us NCCL)
        push = getattr(self, "symm_fused_push", None)
        if push is None:
            push = comm.world * M * N * x.dtype.itemsize <= (1 << 20)
        if push and comm.fits_push(M, N, x.dtype):
            # decode shapes: the GEMM epilogue pushes the partials to every rank through the NVSwitch multicast
            # mapping; what is left of the all-reduce is one barrier + a local rank-ordered sum
            with capi.static_weights():
                if isinstance(self.quant_method, W8A8Int8LinearMethod):
                    q, s = ops.per_token_quant_int8(x2d)
                else:
                    q, s = ops.sglang_per_token_quant_fp8(x2d)
                comm.push_gemm(q, self.weight, s.view(-1), self.weight_scale.view(-1), x.dtype, bias)
                return comm.reduce_gathered(M, N, x.dtype).view(*x.shape[:-1], N)
        # prefill-sized sums: the reduce-scatter half runs inside the GEMM (scatter epilogue over NVLink), the sum is
        # local and the all-gather half is one pass of peer stores -- ahead of NCCL at every rank count and of the
        # two-shot kernel from 32 MB on 8 ranks (profiles/r02w_scatter_{2,8}gpu.log: M = 8192 x 4096, 2 ranks 227 us vs
        # 284 NCCL vs 348 two-shot; 8 ranks 234 vs 318 vs 243)
        fused = getattr(self, "symm_fused_scatter", None)
        if fused is None:
            fused = M * N * x.dtype.itemsize >= ((8 << 20) if comm.world < 8 else (32 << 20))
        if fused and comm.fits_scatter_all_reduce(M, N, x.dtype):
            with capi.static_weights():
                if isinstance(self.quant_method, W8A8Int8LinearMethod):
                    q, s = ops.per_token_quant_int8(x2d)
                else:
                    q, s = ops.sglang_per_token_quant_fp8(x2d)
                comm.scatter_gemm(q, self.weight, s.view(-1), self.weight_scale.view(-1), x.dtype, bias)
                return comm.all_reduce_scattered(M, N, x.dtype).view(*x.shape[:-1], N)
        if not comm.fits(M, N, x.dtype):
            return None
        # prefill-sized messages: the NVSwitch two-shot kernel is ahead of NCCL on 8 ranks (1407 vs 1471 us per decoder
        # layer at M = 8192) and behind it on 2 and 4 (profiles/r02j_bench_2gpu.json, r02t_bench_4gpu.json)
        if (M * N * x.dtype.itemsize > (8 << 20) and comm.world < 8 and not getattr(self, "symm_force", False)):
            return None
        with capi.static_weights():
            if isinstance(self.quant_method, W8A8Int8LinearMethod):
                q, s = ops.per_token_quant_int8(x2d)
            else:
                q, s = ops.sglang_per_token_quant_fp8(x2d)
            part = ops.scaled_mm_out(q, self.weight, s.view(-1), self.weight_scale.view(-1), comm.partial(M, N, x.dtype),
                                     bias)
        return comm.all_reduce(part).view(*x.shape[:-1], N)

    def forward_allreduce_rmsnorm_quant(self, x: torch.Tensor, norm_weight: torch.Tensor, eps: float,
                                        residual: Optional[torch.Tensor] = None, mode: str = "int8",
                                        return_normed: bool = False):
        """Row-parallel layer + all-reduce + fused_add_rmsnorm + the NEXT layer's activation quantizer: what follows
        o_proj / down_proj in a decoder block (RMSNorm.forward_with_allreduce_fusion, python/sglang/srt/layers/
        layernorm.py:300-340 -> GroupCoordinator.fused_allreduce_rmsnorm, parallel_state.py:648-698, then the quantizer
        of qkv_proj / gate_up_proj).  `residual` [M, N] is updated in place to sum + residual.  Returns (q, scales[,
        normed]) as ops.rmsnorm_quant does.  Decode shapes with a SymmAllReduce attached take two launches in total:
        the GEMM pushing its partials to every rank, and b2_allreduce_symm_gathered_rmsnorm_quant."""
        assert self.parallel == "row"
        bias = self.bias if self.tp_rank == 0 else None
        x2d = x.view(-1, x.shape[-1])
        M, N = x2d.shape[0], self.weight.shape[1]
        comm = getattr(self, "symm_comm", None)
        if (self.tp_size > 1 and comm is not None and x.is_cuda and comm.fits_reduce_rmsnorm(M, N, x.dtype)
                and isinstance(self.quant_method, (W8A8Int8LinearMethod, W8A8Fp8LinearMethod))):
            with capi.static_weights():
                if isinstance(self.quant_method, W8A8Int8LinearMethod):
                    q, s = ops.per_token_quant_int8(x2d)
                else:
                    q, s = ops.sglang_per_token_quant_fp8(x2d)
                comm.push_gemm(q, self.weight, s.view(-1), self.weight_scale.view(-1), x.dtype, bias)
                return comm.reduce_gathered_rmsnorm_quant(M, N, x.dtype, norm_weight, eps, residual=residual, mode=mode,
                                                          return_normed=return_normed)
        out = self.forward(x).view(M, N)
        return ops.rmsnorm_quant(out, norm_weight, eps, residual=residual, mode=mode, return_normed=return_normed)

    def forward_reduce_scatter_rmsnorm_quant(self, x: torch.Tensor, norm_weight: torch.Tensor, eps: float,
                                             residual_local: Optional[torch.Tensor] = None, mode: str = "int8"):
        """Prefill form of forward_allreduce_rmsnorm_quant with a sequence-parallel norm: the row-parallel GEMM scatters
        its partials to the row owners, every rank sums, adds ITS rows of the residual (`residual_local`
        [M / tp, N], updated in place; LayerCommunicator._tp_reduce_scatter splits the residual the same way,
        communicator.py:542-560), norms and quantizes its M / tp rows, and the 1-byte codes + scales are all-gathered
        by peer stores.  Returns (q [M, N], scales [M, 1]) for the next column-parallel layer -- views of the
        communicator's buffer when the fused path runs (valid until its next use)."""
        assert self.parallel == "row" and self.tp_size > 1
        bias = self.bias if self.tp_rank == 0 else None
        x2d = x.view(-1, x.shape[-1])
        M, N = x2d.shape[0], self.weight.shape[1]
        comm = getattr(self, "symm_comm", None)
        if (comm is not None and x.is_cuda and mode in ("int8", "fp8") and comm.fits_scatter_rmsnorm_allgather(M, N, x.dtype)
                and isinstance(self.quant_method, (W8A8Int8LinearMethod, W8A8Fp8LinearMethod))):
            with capi.static_weights():
                if isinstance(self.quant_method, W8A8Int8LinearMethod):
                    q, s = ops.per_token_quant_int8(x2d)
                else:
                    q, s = ops.sglang_per_token_quant_fp8(x2d)
                comm.scatter_gemm(q, self.weight, s.view(-1), self.weight_scale.view(-1), x.dtype, bias)
                return comm.reduce_scattered_rmsnorm_quant_allgather(M, N, x.dtype, norm_weight, eps,
                                                                     residual_local=residual_local, mode=mode)
        rows = self.forward_reduce_scatter(x)
        ql, sl = ops.rmsnorm_quant(rows, norm_weight, eps, residual=residual_local, mode=mode)
        q_all = torch.empty(M, N, dtype=ql.dtype, device=x.device)
        s_all = torch.empty(M, 1, dtype=torch.float32, device=x.device)
        dist.all_gather_into_tensor(q_all.view(torch.uint8), ql.view(torch.uint8).contiguous(), group=self.tp_group)
        dist.all_gather_into_tensor(s_all, sl.contiguous(), group=self.tp_group)
        return q_all, s_all

    def forward_reduce_scatter(self, x: torch.Tensor) -> torch.Tensor:
        """Row-parallel layer whose summed output is wanted ROW-SHARDED: returns rows [tp_rank * M/tp, (tp_rank+1) * M/tp)
        of the sum -- RowParallelLinear.forward(skip_all_reduce=True) followed by LayerCommunicator._tp_reduce_scatter
        (python/sglang/srt/layers/linear.py:1467-1490, python/sglang/srt/layers/communicator.py:542-560).  With a
        SymmAllReduce attached and a prefill-sized M the GEMM's scatter epilogue stores every rank's rows straight
        into that rank's gather buffer (b2_gemm_*_scatter_symm) and one barrier + a local rank-ordered fp32 sum
        finish it; otherwise GEMM + reduce_scatter_tensor of the group's backend."""
        assert self.parallel == "row" and self.tp_size > 1
        bias = self.bias if self.tp_rank == 0 else None
        x2d = x.view(-1, x.shape[-1])
        M, N = x2d.shape[0], self.weight.shape[1]
        assert M % self.tp_size == 0, f"Expected total tokens {M} % tp_size {self.tp_size} to be 0"
        comm = getattr(self, "symm_comm", None)
        if (comm is not None and x.is_cuda and comm.fits_scatter(M, N, x.dtype)
                and isinstance(self.quant_method, (W8A8Int8LinearMethod, W8A8Fp8LinearMethod))):
            with capi.static_weights():
                if isinstance(self.quant_method, W8A8Int8LinearMethod):
                    q, s = ops.per_token_quant_int8(x2d)
                else:
                    q, s = ops.sglang_per_token_quant_fp8(x2d)
                comm.scatter_gemm(q, self.weight, s.view(-1), self.weight_scale.view(-1), x.dtype, bias)
                return comm.reduce_scattered(M, N, x.dtype)
        part = self.quant_method.apply(self, x2d, bias)
        out = part.new_empty(M // self.tp_size, N)
        if part.is_cuda:
            dist.reduce_scatter_tensor(out, part, group=self.tp_group)
        else:  # gloo has no reduce_scatter_tensor: all-reduce and keep this rank's rows
            dist.all_reduce(part, group=self.tp_group)
            out.copy_(part[self.tp_rank * (M // self.tp_size):(self.tp_rank + 1) * (M // self.tp_size)])
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        bias = self.bias
        if self.parallel == "row" and self.tp_rank != 0:
            bias = None
        if self.parallel == "row" and self.tp_size > 1:
            out = self._row_parallel_symm(x, bias)
            if out is not None:
                return out
        out = self.quant_method.apply(self, x, bias)
        if self.parallel == "row" and self.tp_size > 1:
            dist.all_reduce(out, group=self.tp_group)  # in place, on the compute stream (NCCL) / gloo on CPU
        return out


# ----------------------------------------------------------- sglang plug-in ----

def install_into_sglang() -> List[str]:
    """Monkey-patch an importable sglang / sgl_kernel so that their W8A8 linear methods call
    this library (SURVEY 7.2 "coexistence").  Returns the names that were patched; with
    neither package importable it returns [] (the torch.ops.sgl_kernel.* registration done
    at import time is then the whole integration)."""
    patched: List[str] = []
    try:
        import sgl_kernel  # type: ignore

        for name in ("int8_scaled_mm", "fp8_scaled_mm", "fp8_blockwise_scaled_mm", "sgl_per_token_quant_fp8"):
            setattr(sgl_kernel, name, getattr(ops, name))
            patched.append(f"sgl_kernel.{name}")
    except Exception:
        pass
    try:
        from sglang.srt.layers.quantization import fp8_kernel, fp8_utils, int8_kernel, w8a8_int8  # type: ignore

        int8_kernel.per_token_quant_int8 = ops.per_token_quant_int8
        w8a8_int8.per_token_quant_int8 = ops.per_token_quant_int8
        w8a8_int8.int8_scaled_mm = ops.int8_scaled_mm
        fp8_kernel.sglang_per_token_quant_fp8 = ops.sglang_per_token_quant_fp8
        fp8_utils.sglang_per_token_quant_fp8 = ops.sglang_per_token_quant_fp8
        fp8_utils.fp8_scaled_mm = ops.fp8_scaled_mm
        fp8_utils.fp8_blockwise_scaled_mm = ops.fp8_blockwise_scaled_mm
        patched += ["sglang.srt.layers.quantization.{int8_kernel,w8a8_int8,fp8_kernel,fp8_utils}"]
    except Exception:
        pass
    return patched
