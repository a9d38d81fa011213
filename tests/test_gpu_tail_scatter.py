"""GPU tests (-m gpu, ONE GPU) of two modes of the persistent GEMM added in round 2:

* tail split -- the ragged last wave of 256-wide tiles issued as half-width work items -- must not change a bit
  (int32 accumulation is exact; the fp8 K order is untouched), including ragged N / M and whole problems that are
  nothing but tail;
* scatter epilogue (b2_gemm_*_scatter_symm, the prefill reduce-scatter of row-parallel layers): exercised here with
  every "rank's" gather buffer on the same device -- what a rank stores into slot `rank` of every owner's buffer must be
  the owner's rows of the plain GEMM, for both sets; the cross-GPU sum is covered by tests/test_multi_gpu.py.
"""
from __future__ import annotations

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

import sglang_diffusion_b200 as b2  # noqa: E402
from oracle import restate  # noqa: E402  (the checker)

capi = b2.capi
DEV = "cuda:0"
NO_TAIL = 8 << 8  # developer flag 8 in the forced-config word: tail split off


def _problem(kind, M, N, K, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    if kind == "int8":
        A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g, device=DEV)
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=DEV)
    else:
        A = torch.randn(M, K, generator=g, device=DEV).to(torch.float8_e4m3fn)
        W = torch.randn(N, K, generator=g, device=DEV).to(torch.float8_e4m3fn)
    As = torch.rand(M, 1, generator=g, device=DEV) * 1e-2 + 1e-3
    Bs = torch.rand(N, 1, generator=g, device=DEV) * 1e-2 + 1e-3
    bias = torch.randn(N, generator=g, device=DEV).bfloat16()
    return A, W, As, Bs, bias


TAIL_SHAPES = [
    (2048, 5120, 512),    # 160 tiles on 74 CTA pairs: 2 waves + 12 -> 24 half tiles
    (1280, 1032, 384),    # 25 tiles, all of them tail; last column block holds 8 columns
    (4000, 5120, 256),    # ragged M
    (1100, 2600, 384),    # ragged M and N, N % 64 != 0
    (4096, 5120, 1024),   # 320 tiles: 4 waves + 24
    (640, 520, 128),      # a single K block
]


@pytest.mark.parametrize("kind", ["int8", "fp8"])
@pytest.mark.parametrize("M,N,K", TAIL_SHAPES)
@pytest.mark.parametrize("cfg", [0, 4, 5])
def test_tail_split_is_bit_identical(kind, M, N, K, cfg):
    A, W, As, Bs, bias = _problem(kind, M, N, K, M + N + K)
    op = b2.int8_scaled_mm if kind == "int8" else b2.fp8_scaled_mm
    outs = {}
    for name, word in (("split", cfg), ("nosplit", cfg | NO_TAIL)):
        capi.set_gemm_config(word)
        try:
            outs[name] = op(A, W.t(), As, Bs, torch.bfloat16, bias)
        finally:
            capi.set_gemm_config(0)
    torch.cuda.synchronize()
    assert torch.equal(outs["split"].view(torch.int16), outs["nosplit"].view(torch.int16))
    if kind == "int8":  # and both equal the oracle on a sample of rows (bit-exact contract of the int8 path)
        rows = torch.linspace(0, M - 1, 24).long().unique()
        a_u8 = (A[rows.to(DEV)].cpu().to(torch.int16) + 128).to(torch.uint8)
        want = restate.int8_scaled_mm_cpu(a_u8, W.cpu(), As[rows.to(DEV)].cpu(), Bs.cpu(), bias.cpu())
        got = outs["split"][rows.to(DEV)].cpu()
        assert torch.equal(got.view(torch.int16), want.view(torch.int16))


def _scatter(kind, A, W, As, Bs, bias, bufs, pad, set_stride, rank, world, out_dtype):
    M, K = A.shape
    N = W.shape[0]
    fn = capi.lib.b2_gemm_s8_scatter_symm if kind == "int8" else capi.lib.b2_gemm_e4m3_scatter_symm
    peers = (ctypes.c_void_p * world)(*[b.data_ptr() for b in bufs])
    od = capi.DTYPE_BF16 if out_dtype == torch.bfloat16 else capi.DTYPE_F16
    capi.check(fn(A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0), As.data_ptr(), Bs.data_ptr(),
                  None if bias is None else bias.data_ptr(), capi.DTYPE_BF16, peers, pad.data_ptr(), set_stride, rank,
                  world, od, M, N, K, torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("kind", ["int8", "fp8"])
@pytest.mark.parametrize("world,M,N,K", [(2, 1024, 1024, 512), (4, 1280, 1032, 384), (8, 2048, 4096, 256),
                                         (8, 256, 520, 128), (3, 96 * 3, 768, 1024)])
@pytest.mark.parametrize("counter", [0, 1, 6])
def test_scatter_epilogue_fills_the_owners_slots(kind, world, M, N, K, counter):
    rows = M // world
    slot = rows * N * 2
    set_stride = (world * slot + 4096 + 15) // 16 * 16
    bufs = [torch.full((2 * set_stride,), 0x5A, dtype=torch.uint8, device=DEV) for _ in range(world)]
    pad = torch.zeros(2048, dtype=torch.int32, device=DEV)
    word = 1544  # kGatherCounterWord of csrc/allreduce_symm.cu (include/b200_w8a8.h documents it)
    pad[word] = counter
    plain = []
    op = b2.int8_scaled_mm if kind == "int8" else b2.fp8_scaled_mm
    for r in range(world):
        A, W, As, Bs, bias = _problem(kind, M, N, K, 17 * r + M + N)
        b = bias if r == 0 else None
        plain.append(op(A, W.t(), As, Bs, torch.bfloat16, b))
        _scatter(kind, A, W, As.view(-1), Bs.view(-1), b, bufs, pad, set_stride, r, world, torch.bfloat16)
    torch.cuda.synchronize()
    s = counter & 1
    for owner in range(world):
        cur = bufs[owner][s * set_stride: s * set_stride + world * slot].view(torch.bfloat16).view(world, rows, N)
        for r in range(world):
            want = plain[r][owner * rows:(owner + 1) * rows]
            assert torch.equal(cur[r].view(torch.int16), want.view(torch.int16)), (owner, r)
        # nothing outside the slots of the set in use was touched
        other = bufs[owner][(1 - s) * set_stride:(1 - s) * set_stride + world * slot]
        assert bool((other == 0x5A).all())
        gap = bufs[owner][s * set_stride + world * slot: s * set_stride + world * slot + 4096] if s == 0 else \
            bufs[owner][s * set_stride + world * slot:]
        assert bool((gap == 0x5A).all())
    assert int(pad[word].item()) == counter  # the GEMM only reads the counter


def test_scatter_rejects_rows_that_do_not_split_into_store_boxes():
    A, W, As, Bs, _ = _problem("int8", 100, 256, 128, 3)
    bufs = [torch.zeros(1 << 20, dtype=torch.uint8, device=DEV) for _ in range(2)]
    pad = torch.zeros(2048, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError):
        _scatter("int8", A, W, As.view(-1), Bs.view(-1), None, bufs, pad, 0, 0, 2, torch.bfloat16)
