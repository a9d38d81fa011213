"""Host-buffer entry to the W8A8-int8 linear: the end-to-end call for a caller whose activations
live in HOST memory (what the reference's CPU op int8_scaled_mm_with_quant,
sgl-kernel/csrc/cpu/gemm_int8.cpp:481-547, is given).  Weights, scales and bias stay resident in HBM
(they are model state, loaded once -- process_weights_after_loading); every call moves the step's
activations host->device and the step's outputs device->host.

The copy engines and the SMs are overlapped: rows are cut into chunks and each chunk flows through
three streams -- H2D copy, quantize+GEMM (one C-ABI call, b2_linear_w8a8_int8), D2H copy -- with
double-buffered device staging, so the call runs at PCIe speed rather than copy + compute + copy.
No arithmetic happens here; this module only moves bytes and enqueues library calls.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import capi


class HostW8A8Int8Linear:
    def __init__(self, weight: torch.Tensor, weight_scale: torch.Tensor, bias: Optional[torch.Tensor] = None,
                 device: Optional[torch.device] = None, chunk_rows: int = 2048,
                 act_dtype: torch.dtype = torch.bfloat16):
        """weight: int8 [N,K] (checkpoint layout), weight_scale: fp32 [N] / [N,1], bias: [N] or None."""
        assert weight.dtype == torch.int8 and weight.dim() == 2
        self.device = torch.device(device if device is not None else "cuda")
        self.N, self.K = weight.shape
        self.W = weight.to(self.device).contiguous()
        self.Bs = weight_scale.reshape(-1).float().to(self.device).contiguous()
        self.bias = None if bias is None else bias.to(self.device).contiguous()
        self.act_dtype = act_dtype
        self.chunk = int(chunk_rows)
        self._xbuf = [torch.empty(self.chunk, self.K, dtype=act_dtype, device=self.device) for _ in range(2)]
        self._obuf = [torch.empty(self.chunk, self.N, dtype=act_dtype, device=self.device) for _ in range(2)]
        self._ws_bytes = int(capi.lib.b2_linear_w8a8_int8_workspace(self.chunk, self.K))
        self._ws = [torch.empty(self._ws_bytes, dtype=torch.uint8, device=self.device) for _ in range(2)]
        self.s_in, self.s_comp, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _dt(self, t: torch.dtype) -> int:
        return {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}[t]

    def __call__(self, x_host: torch.Tensor, out_host: torch.Tensor) -> torch.Tensor:
        """x_host [M,K] and out_host [M,N]: CPU tensors (pinned for full PCIe speed).  Returns out_host
        after everything has landed (the call synchronises its own streams, nothing else)."""
        self.submit(x_host, out_host)
        return self.wait()

    def wait(self) -> torch.Tensor:
        """Block until the outputs of the last submit() are in host memory."""
        self.s_out.synchronize()
        return self._last_out

    def submit(self, x_host: torch.Tensor, out_host: torch.Tensor) -> None:
        """Enqueue the whole call (copies in, quantize + GEMM, copies out) on this layer's own streams and return
        at once.  Independent layers submitted back to back share the link in both directions -- one layer's
        uploads run under another's downloads -- which a caller with several linears over host-resident
        activations (QKV / gate_up of one token batch) gets for free; wait() then collects each.  One submit may be
        in flight per layer object (the staging buffers are per layer): wait() before the next submit()."""
        assert x_host.device.type == "cpu" and out_host.device.type == "cpu"
        assert x_host.dtype == self.act_dtype and out_host.dtype == self.act_dtype
        M = x_host.shape[0]
        assert x_host.shape == (M, self.K) and out_host.shape == (M, self.N)
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_in, self.s_comp, self.s_out):
            s.wait_stream(cur)
        comp_done, out_done = [], []
        nchunks = (M + self.chunk - 1) // self.chunk
        for i in range(nchunks):
            r0 = i * self.chunk
            r1 = min(M, r0 + self.chunk)
            rows = r1 - r0
            b = i & 1
            with torch.cuda.stream(self.s_in):
                if i >= 2:
                    self.s_in.wait_event(comp_done[i - 2])      # staging buffer b is free again
                self._xbuf[b][:rows].copy_(x_host[r0:r1], non_blocking=True)
                in_done = torch.cuda.Event()
                in_done.record(self.s_in)
            with torch.cuda.stream(self.s_comp):
                self.s_comp.wait_event(in_done)
                if i >= 2:
                    self.s_comp.wait_event(out_done[i - 2])     # output buffer b has been drained
                capi.check(capi.lib.b2_linear_w8a8_int8(
                    self._xbuf[b].data_ptr(), self._dt(self.act_dtype), self.K, self.W.data_ptr(), self.K,
                    self.Bs.data_ptr(), None if self.bias is None else self.bias.data_ptr(),
                    capi.DTYPE_BF16 if self.bias is None else self._dt(self.bias.dtype),
                    self._obuf[b].data_ptr(), self.N, rows, self.N, self.K, self._ws[b].data_ptr(), self._ws_bytes,
                    self.s_comp.cuda_stream))
                ev = torch.cuda.Event()
                ev.record(self.s_comp)
                comp_done.append(ev)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(comp_done[i])
                out_host[r0:r1].copy_(self._obuf[b][:rows], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.s_out)
                out_done.append(ev)
        self.h2d_bytes += x_host.numel() * x_host.element_size()
        self.d2h_bytes += out_host.numel() * out_host.element_size()
        self._last_out = out_host
