"""Peer-memory all-reduce for the row-parallel W8A8 linear layers (decode-sized messages).

The reference sums the partial outputs of a RowParallelLinear with `tensor_model_parallel_all_reduce`
(python/sglang/srt/layers/linear.py:1486-1490), which picks its custom all-reduce kernel for small messages
(sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:204-260, python/sglang/srt/distributed/device_communicators/
custom_all_reduce.py:328-346) and NCCL otherwise.  Here the small-message path is `b2_allreduce_symm`
(csrc/allreduce_symm.cu): the GEMM writes its partial output into a symmetric buffer and one kernel reduces it
through the NVSwitch (multimem.ld_reduce / multimem.st) or, without a multicast mapping, with ordered peer loads.

PyTorch is plumbing only: `torch.distributed._symmetric_memory` allocates the buffer and exchanges the peer /
multicast mappings; the data path is this library's kernel.  Larger messages keep NCCL.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import capi

_DT = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16}


class SymmAllReduce:
    """One symmetric buffer of `max_bytes` per rank of `group`; every rank must construct it collectively."""

    def __init__(self, group=None, max_bytes: int = 8 << 20, device: Optional[torch.device] = None,
                 full_bytes: int = 0):
        import torch.distributed._symmetric_memory as symm_mem  # raises where the build has no symmetric memory

        group = group if group is not None else dist.group.WORLD
        self.group = group
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.max_bytes = int(max_bytes)
        self.buf = symm_mem.empty(self.max_bytes, dtype=torch.uint8, device=self.device)
        self.handle = symm_mem.rendezvous(self.buf, group.group_name)
        self.rank, self.world = int(self.handle.rank), int(self.handle.world_size)
        if not 2 <= self.world <= 8:
            raise ValueError("SymmAllReduce serves 2..8 ranks of one NVSwitch domain")
        self.multicast_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.bufs_dev = int(self.handle.buffer_ptrs_dev)
        self.pads_dev = int(self.handle.signal_pad_ptrs_dev)
        # signal pad layout of csrc/allreduce_symm.cu: words [0, world * blocks) flags, [1024, 1024 + blocks) counters,
        # word 1536 the time-out record.  The pad of this handle is owned by b2_allreduce_symm (barrier numbers only
        # grow and are never reset): do not call handle.barrier() / put_signal() / wait_signal() on it.
        self.max_blocks = min(128, 1024 // self.world)
        # gather form (push / scatter epilogues): two sets, a fixed half buffer apart whatever the size of the call
        self.set_stride = (self.max_bytes // 2) // 16 * 16
        import ctypes
        self._peer_bufs_host = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        # second symmetric buffer: the all-gather half of the fused prefill all-reduce lands here (multicast stores)
        self.full_buf = None
        self.full_bufs_dev = 0
        if full_bytes > 0:
            self.full_buf = symm_mem.empty(int(full_bytes), dtype=torch.uint8, device=self.device)
            fh = symm_mem.rendezvous(self.full_buf, group.group_name)
            self._full_handle = fh
            self.full_bufs_dev = int(fh.buffer_ptrs_dev)
        need = 4 * (1536 + 1)
        if int(self.handle.signal_pad_size) < need:
            raise ValueError(f"signal pad of {self.handle.signal_pad_size} bytes is smaller than the {need} b2_allreduce_symm needs")

    @staticmethod
    def set_timeout_ms(ms: int) -> None:
        """Bound every wait on a peer (b2_set_allreduce_timeout_ms; 0 = wait for ever, the default)."""
        capi.lib.b2_set_allreduce_timeout_ms(int(ms))

    def timed_out(self) -> int:
        """Barrier number at which a wait of this rank gave up (0 = never).  Synchronises the device."""
        pad = self.handle.get_signal_pad(self.rank, (1537,), torch.uint32)
        return int(pad[1536].item())

    # ---- GEMM-epilogue-fused form (b2_gemm_*_push_symm + b2_allreduce_symm_gathered) ----

    def fits_push(self, M: int, N: int, dtype: torch.dtype) -> bool:
        """The decode-shape GEMM can push its partials through the multicast mapping: needs the NVSwitch mapping, M <=
        128 and room for two sets of `world` slots."""
        return (self.multicast_ptr != 0 and dtype in _DT and 0 < M <= 128 and (M * N) % 8 == 0
                and self.world * M * N * dtype.itemsize <= self.set_stride)

    def push_gemm(self, q: torch.Tensor, weight: torch.Tensor, a_scale: torch.Tensor, w_scale: torch.Tensor,
                  out_dtype: torch.dtype, bias: Optional[torch.Tensor] = None) -> None:
        """int8 / e4m3 scaled GEMM (chosen by q.dtype; weight is [K,N] with stride(0) == 1) whose epilogue stores this
        rank's partial output into slot `rank` of every rank's gather buffer."""
        M, K = q.shape
        N = weight.shape[1]
        assert weight.shape[0] == K and weight.stride(0) == 1 and q.stride(1) == 1
        fn = capi.lib.b2_gemm_s8_push_symm if q.dtype == torch.int8 else capi.lib.b2_gemm_e4m3_push_symm
        bdt = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}
        with torch.cuda.device(q.device):
            capi.check(fn(q.data_ptr(), q.stride(0) if M > 1 else K, weight.data_ptr(), weight.stride(1) if N > 1 else K,
                          a_scale.data_ptr(), w_scale.data_ptr(), None if bias is None else bias.data_ptr(),
                          bdt[bias.dtype] if bias is not None else capi.DTYPE_F32, self.multicast_ptr,
                          int(self.handle.signal_pad_ptrs[self.rank]), self.set_stride, self.rank, self.world,
                          _DT[out_dtype], M, N, K, torch.cuda.current_stream(q.device).cuda_stream))

    def reduce_gathered(self, M: int, N: int, dtype: torch.dtype) -> torch.Tensor:
        """Barrier + rank-ordered fp32 sum of the `world` pushed slots (local memory) -> fresh [M,N] tensor."""
        out = torch.empty(M, N, dtype=dtype, device=self.device)
        with torch.cuda.device(self.device):
            capi.check(capi.lib.b2_allreduce_symm_gathered(self.buf.data_ptr(), self.pads_dev, out.data_ptr(), M * N,
                                                           self.set_stride, _DT[dtype], self.rank, self.world,
                                                           self.max_blocks,
                                                           torch.cuda.current_stream(self.device).cuda_stream))
        return out

    def fits_reduce_rmsnorm(self, M: int, N: int, dtype: torch.dtype) -> bool:
        return self.fits_push(M, N, dtype) and N <= 32768

    def reduce_gathered_rmsnorm_quant(self, M: int, N: int, dtype: torch.dtype, weight: torch.Tensor, eps: float,
                                      residual: Optional[torch.Tensor] = None, mode: str = "int8",
                                      return_normed: bool = False):
        """After push_gemm: the all-reduce's barrier + local sum, the residual add (in place), RMSNorm and the next
        layer's activation quantizer in ONE launch (b2_allreduce_symm_gathered_rmsnorm_quant; the reference's
        fused_allreduce_rmsnorm hook, parallel_state.py:648-698, plus the quantizer that follows it).
        mode "int8" | "fp8" | "fp8_group128".  Returns (q, scales[, normed]) like ops.rmsnorm_quant."""
        from . import ops

        qm = {"int8": 0, "fp8": 1, "fp8_group128": 2}[mode]
        assert weight.is_contiguous() and weight.dtype == dtype and weight.numel() == N
        if residual is not None:
            assert residual.is_contiguous() and residual.shape == (M, N) and residual.dtype == dtype
        q = torch.empty(M, N, dtype=torch.int8 if qm == 0 else ops.fp8_dtype, device=self.device)
        if qm == 2:
            s = ops.create_per_token_group_quant_fp8_output_scale((M, N), self.device, 128, True)
            srow, sgrp = s.stride(0), s.stride(1)
        else:
            s = torch.empty(M, 1, dtype=torch.float32, device=self.device)
            srow, sgrp = 1, 0
        y = torch.empty(M, N, dtype=dtype, device=self.device) if return_normed else None
        with torch.cuda.device(self.device):
            capi.check(capi.lib.b2_allreduce_symm_gathered_rmsnorm_quant(
                self.buf.data_ptr(), self.pads_dev, None if residual is None else residual.data_ptr(), N,
                weight.data_ptr(), float(eps), qm, q.data_ptr(), N, s.data_ptr(), srow, sgrp,
                None if y is None else y.data_ptr(), N, M, N, self.set_stride, _DT[dtype], self.rank, self.world,
                self.max_blocks, torch.cuda.current_stream(self.device).cuda_stream))
        return (q, s, y) if return_normed else (q, s)

    # ---- prefill: scatter epilogue + gathered sum = fused reduce-scatter (b2_gemm_*_scatter_symm) ----

    def fits_scatter(self, M: int, N: int, dtype: torch.dtype) -> bool:
        """The persistent GEMM can store every rank's rows straight into that rank's gather buffer."""
        return (dtype in _DT and M > 0 and M % (32 * self.world) == 0 and N % 8 == 0
                and M * N * dtype.itemsize <= self.set_stride)

    def scatter_gemm(self, q: torch.Tensor, weight: torch.Tensor, a_scale: torch.Tensor, w_scale: torch.Tensor,
                     out_dtype: torch.dtype, bias: Optional[torch.Tensor] = None) -> None:
        """int8 / e4m3 scaled GEMM whose epilogue stores rows [r*M/world, (r+1)*M/world) of this rank's partial output
        into slot `rank` of rank r's gather buffer (TMA stores to the peer mappings)."""
        M, K = q.shape
        N = weight.shape[1]
        assert weight.shape[0] == K and weight.stride(0) == 1 and q.stride(1) == 1
        fn = capi.lib.b2_gemm_s8_scatter_symm if q.dtype == torch.int8 else capi.lib.b2_gemm_e4m3_scatter_symm
        bdt = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}
        with torch.cuda.device(q.device):
            capi.check(fn(q.data_ptr(), q.stride(0), weight.data_ptr(), weight.stride(1) if N > 1 else K,
                          a_scale.data_ptr(), w_scale.data_ptr(), None if bias is None else bias.data_ptr(),
                          bdt[bias.dtype] if bias is not None else capi.DTYPE_F32, self._peer_bufs_host,
                          int(self.handle.signal_pad_ptrs[self.rank]), self.set_stride, self.rank, self.world,
                          _DT[out_dtype], M, N, K, torch.cuda.current_stream(q.device).cuda_stream))

    def fits_scatter_all_reduce(self, M: int, N: int, dtype: torch.dtype) -> bool:
        return (self.fits_scatter(M, N, dtype) and self.full_bufs_dev != 0
                and M * N * dtype.itemsize <= self.full_buf.numel())

    def all_reduce_scattered(self, M: int, N: int, dtype: torch.dtype, inplace: bool = False) -> torch.Tensor:
        """After scatter_gemm: barrier, rank-ordered fp32 sum of this rank's rows, peer stores of the result into every
        rank's `full_buf`, barrier -> the whole [M, N] sum (a fresh tensor, or -- inplace=True -- a view of full_buf that
        stays valid until the next call using it).  b2_allreduce_symm_gathered_bcast."""
        out = None if inplace else torch.empty(M, N, dtype=dtype, device=self.device)
        with torch.cuda.device(self.device):
            capi.check(capi.lib.b2_allreduce_symm_gathered_bcast(
                self.buf.data_ptr(), self.pads_dev, self.full_bufs_dev, None if inplace else out.data_ptr(), (M // self.world) * N, self.set_stride, _DT[dtype], self.rank,
                self.world, self.max_blocks, torch.cuda.current_stream(self.device).cuda_stream))
        if inplace:
            return self.full_buf[:M * N * dtype.itemsize].view(dtype).view(M, N)
        return out

    def fits_scatter_rmsnorm_allgather(self, M: int, N: int, dtype: torch.dtype) -> bool:
        return (self.fits_scatter(M, N, dtype) and self.full_bufs_dev != 0 and N % 16 == 0 and N <= 32768
                and self._act_scale_offset(M, N) + 4 * M <= self.full_buf.numel())

    @staticmethod
    def _act_scale_offset(M: int, N: int) -> int:
        return (M * N + 255) // 256 * 256

    def reduce_scattered_rmsnorm_quant_allgather(self, M: int, N: int, dtype: torch.dtype, weight: torch.Tensor,
                                                 eps: float, residual_local: Optional[torch.Tensor] = None,
                                                 mode: str = "int8"):
        """After scatter_gemm: sum of this rank's M/world rows + residual add (its rows, in place) + RMSNorm + per-token
        quantizer, codes and scales all-gathered into every rank's `full_buf` by peer stores
        (b2_reduce_scatter_symm_rmsnorm_quant_allgather).  Returns (q [M, N], scales [M, 1]) -- VIEWS of full_buf,
        valid until the next call that uses it: the input of the next column-parallel GEMM."""
        from . import ops

        qm = {"int8": 0, "fp8": 1}[mode]
        rows = M // self.world
        assert weight.is_contiguous() and weight.dtype == dtype and weight.numel() == N
        if residual_local is not None:
            assert residual_local.is_contiguous() and residual_local.shape == (rows, N) and residual_local.dtype == dtype
        off = self._act_scale_offset(M, N)
        with torch.cuda.device(self.device):
            capi.check(capi.lib.b2_reduce_scatter_symm_rmsnorm_quant_allgather(
                self.buf.data_ptr(), self.pads_dev, None if residual_local is None else residual_local.data_ptr(), N,
                weight.data_ptr(), float(eps), qm, self.full_bufs_dev, off, None, N, M, N, self.set_stride, _DT[dtype],
                self.rank, self.world, self.max_blocks, torch.cuda.current_stream(self.device).cuda_stream))
        q = self.full_buf[:M * N].view(torch.int8 if qm == 0 else ops.fp8_dtype).view(M, N)
        s = self.full_buf[off:off + 4 * M].view(torch.float32).view(M, 1)
        return q, s

    def reduce_scattered(self, M: int, N: int, dtype: torch.dtype) -> torch.Tensor:
        """Barrier + rank-ordered fp32 sum of the `world` slots of this rank's rows -> fresh [M / world, N] tensor."""
        return self.reduce_gathered(M // self.world, N, dtype)

    def partial(self, M: int, N: int, dtype: torch.dtype) -> torch.Tensor:
        """[M,N] view of the local symmetric buffer: the GEMM's output operand."""
        nbytes = M * N * dtype.itemsize
        if nbytes > self.max_bytes:
            raise ValueError(f"message of {nbytes} bytes exceeds the symmetric buffer ({self.max_bytes})")
        return self.buf[:nbytes].view(dtype).view(M, N)

    def fits(self, M: int, N: int, dtype: torch.dtype) -> bool:
        return dtype in _DT and M * N * dtype.itemsize <= self.max_bytes and (M * N) % 8 == 0 and M * N > 0

    def all_reduce(self, partial: torch.Tensor, algo: int = 0, inplace: bool = False) -> torch.Tensor:
        """Sum of `partial` (a tensor returned by .partial(), already written by this rank's GEMM on the current
        stream) over the ranks; returns a fresh private tensor, or -- inplace=True, multicast only -- `partial`
        itself holding the sums (valid until the next call that uses this buffer).  Collective: every rank calls it."""
        out = None if inplace else torch.empty(partial.shape, dtype=partial.dtype, device=partial.device)
        with torch.cuda.device(partial.device):  # the library launches on the current device
            stream = torch.cuda.current_stream(partial.device).cuda_stream
            capi.check(capi.lib.b2_allreduce_symm(self.bufs_dev, self.pads_dev, self.multicast_ptr or None,
                                                  None if inplace else out.data_ptr(), partial.numel(),
                                                  _DT[partial.dtype], self.rank, self.world, algo, self.max_blocks,
                                                  stream))
        return partial if inplace else out
