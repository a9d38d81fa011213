"""Out-of-bounds WRITE detection without compute-sanitizer (closed on this pool, profiles/r02b_sanitizer_unavailable.log):
every output of every kernel family is carved out of the middle of a larger buffer pre-filled with a canary pattern;
after the call the bytes in front of and behind the output (and, for strided outputs, the gaps between rows) must be
untouched.  Shapes are ragged on purpose (rows / columns that do not fill the last tile, vector tail, box or chunk)."""
from __future__ import annotations

import pytest
import torch

import sglang_diffusion_b200 as b2
from sglang_diffusion_b200 import capi, moe, ops
from conftest import mk_fp8_block_weight, mk_fp8_weight, mk_int8_problem

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
PAD = 4096  # canary bytes on each side
CANARY = 0xA5


class Guarded:
    """A [rows, cols] tensor of `dtype` with row stride ld >= cols inside a canary-filled byte buffer."""

    def __init__(self, rows, cols, dtype, ld=None):
        self.esize = torch.empty(0, dtype=dtype).element_size()
        self.rows, self.cols, self.ld = rows, cols, ld or cols
        nbytes = (max(rows, 1) - 1) * self.ld * self.esize + cols * self.esize if rows else 0
        self.nbytes = nbytes
        self.raw = torch.full((PAD + nbytes + PAD + 64,), CANARY, dtype=torch.uint8, device=DEV)
        off = PAD + (-(self.raw.data_ptr() + PAD)) % 16  # 16-byte aligned start
        self.off = off
        self.t = torch.as_strided(self.raw[off: off + nbytes].view(dtype) if nbytes else self.raw[:0].view(dtype),
                                  (rows, cols), (self.ld, 1))

    def check(self, what):
        torch.cuda.synchronize()
        r = self.raw.cpu()
        assert bool((r[: self.off] == CANARY).all()), f"{what}: bytes in front of the output were written"
        assert bool((r[self.off + self.nbytes:] == CANARY).all()), f"{what}: bytes behind the output were written"
        if self.ld > self.cols and self.rows > 1:
            body = r[self.off: self.off + (self.rows - 1) * self.ld * self.esize].view(self.rows - 1, self.ld * self.esize)
            assert bool((body[:, self.cols * self.esize:] == CANARY).all()), f"{what}: row gaps were written"


def _st():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("M,K", [(1, 32), (3, 544), (17, 4096), (130, 1368), (5, 14336)])
def test_row_quantizers_stay_inside_their_outputs(M, K):
    x = (torch.randn(M, K, device=DEV) / 10).bfloat16()
    for fn, qdt in ((capi.lib.b2_quant_int8_rows, torch.int8), (capi.lib.b2_quant_fp8_rows, torch.uint8)):
        q, s = Guarded(M, K, qdt, ld=K + 32), Guarded(1, M, torch.float32)
        capi.check(fn(x.data_ptr(), capi.DTYPE_BF16, K, q.t.data_ptr(), K + 32, s.t.data_ptr(), M, K, _st()))
        q.check("row quantizer codes")
        s.check("row quantizer scales")


@pytest.mark.parametrize("M,K", [(1, 128), (5, 640), (33, 1536), (130, 7168)])
def test_group_and_fused_quantizers_stay_inside_their_outputs(M, K):
    x = (torch.randn(M, K, device=DEV) / 10).bfloat16()
    G = K // 128
    q, s = Guarded(M, K, torch.uint8), Guarded(G, M, torch.float32)  # column-major scales: [G, M] storage
    capi.check(capi.lib.b2_quant_fp8_group128(x.data_ptr(), capi.DTYPE_BF16, K, q.t.data_ptr(), K, s.t.data_ptr(), 1, M,
                                              M, K, 1e-10, -448.0, 448.0, _st()))
    q.check("group quantizer codes"); s.check("group quantizer scales")
    q, s = Guarded(M, K, torch.uint8), Guarded(G, M, torch.float32)
    capi.check(capi.lib.b2_quant_fp8_group128_v2(x.data_ptr(), capi.DTYPE_BF16, K, q.t.data_ptr(), K, s.t.data_ptr(), 1, M,
                                                 M, K, _st()))
    q.check("v2 group quantizer codes"); s.check("v2 group quantizer scales")
    x2 = torch.randn(M, 2 * K, device=DEV).bfloat16()
    w = torch.ones(K, device=DEV).bfloat16()
    for mode in (0, 1, 3):
        q, y = Guarded(M, K, torch.uint8), Guarded(M, K, torch.bfloat16)
        s = Guarded(G, M, torch.float32) if mode == 3 else Guarded(1, M, torch.float32)
        capi.check(capi.lib.b2_silu_mul_quant(x2.data_ptr(), capi.DTYPE_BF16, 2 * K, mode, q.t.data_ptr(), K, s.t.data_ptr(),
                                              1, M if mode == 3 else 0, 0, y.t.data_ptr(), K, None, 0, M, K, _st()))
        q.check(f"silu quant mode {mode} codes"); s.check(f"silu quant mode {mode} scales"); y.check("silu act rows")
    for mode in (0, 1, 2):
        q, y, res = Guarded(M, K, torch.uint8), Guarded(M, K, torch.bfloat16), Guarded(M, K, torch.bfloat16)
        res.t.copy_(torch.randn(M, K, device=DEV).bfloat16())
        s = Guarded(G, M, torch.float32) if mode == 2 else Guarded(1, M, torch.float32)
        capi.check(capi.lib.b2_rmsnorm_quant(x.data_ptr(), capi.DTYPE_BF16, K, res.t.data_ptr(), K, w.data_ptr(), 1e-6, mode,
                                             q.t.data_ptr(), K, s.t.data_ptr(), 1, M if mode == 2 else 0, y.t.data_ptr(), K,
                                             M, K, _st()))
        q.check(f"rmsnorm quant mode {mode} codes"); s.check("rmsnorm quant scales"); y.check("normed rows")
        res.check("residual rows")
    q, s1 = Guarded(1, M * K, torch.uint8), Guarded(1, 1, torch.float32)
    s1.t.zero_()
    capi.check(capi.lib.b2_quant_fp8_tensor(x.data_ptr(), capi.DTYPE_BF16, q.t.data_ptr(), s1.t.data_ptr(), None, M * K, M, 0,
                                            _st()))
    q.check("per-tensor codes"); s1.check("per-tensor scale")


@pytest.mark.parametrize("cfg", [0, 1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("M,N,K", [(1, 8, 16), (5, 264, 544), (130, 1000, 1168), (300, 520, 4096)])
def test_gemm_outputs_stay_inside_their_rows(cfg, M, N, K):
    """Persistent kernel (six tile configs, TMA stores clipped by the tensor map) and the planner's pick (split-K for the
    small shapes: register stores) with a row stride larger than N."""
    x, W, Bs, b = mk_int8_problem(M, N, K, seed=M + N)
    q, s = ops.per_token_quant_int8(x.to(DEV))
    Wd, Bsd, bd = W.to(DEV), Bs.to(DEV), b.to(DEV)
    D = Guarded(M, N, torch.bfloat16, ld=N + 24)
    capi.set_gemm_config(cfg)
    try:
        capi.check(capi.lib.b2_gemm_s8(q.data_ptr(), K, Wd.data_ptr(), K, s.data_ptr(), Bsd.data_ptr(), bd.data_ptr(),
                                       capi.DTYPE_BF16, D.t.data_ptr(), N + 24, capi.DTYPE_BF16, M, N, K, _st()))
    finally:
        capi.set_gemm_config(0)
    D.check(f"int8 GEMM cfg {cfg} ({capi.last_gemm_config()})")
    ref = ops.int8_scaled_mm(q, Wd.t(), s, Bsd.view(-1, 1), torch.bfloat16, bd)
    assert torch.equal(D.t.contiguous().view(torch.int16), ref.view(torch.int16))


@pytest.mark.parametrize("cfg", [0, 1, 2, 7, 8, 25, 41])
@pytest.mark.parametrize("M,N,K", [(1, 128, 128), (37, 384, 640), (130, 1536, 2048)])
def test_blockwise_gemm_outputs_stay_inside(cfg, M, N, K):
    x = (torch.randn(M, K) / 10).bfloat16().to(DEV)
    Wq, sfb = mk_fp8_block_weight(N, K)
    q, sfa = ops.per_token_group_quant_fp8(x, 128, column_major_scales=True)
    Wd, sfbd = Wq.to(DEV), sfb.to(DEV)
    D = Guarded(M, N, torch.bfloat16, ld=N + 8)
    capi.set_gemm_config(cfg)
    try:
        st = capi.lib.b2_gemm_e4m3_block128(q.data_ptr(), K, Wd.data_ptr(), K, sfa.data_ptr(), sfa.stride(0), sfa.stride(1),
                                            sfbd.data_ptr(), sfbd.stride(0), sfbd.stride(1), None, 0, D.t.data_ptr(), N + 8,
                                            capi.DTYPE_BF16, M, N, K, _st())
    finally:
        capi.set_gemm_config(0)
    if st != 0 and cfg in (25, 41):  # a forced split that does not fit this K is refused, not launched
        assert "does not fit" in capi.last_error()
        return
    capi.check(st)
    D.check(f"block-wise GEMM cfg {cfg} ({capi.last_gemm_config()})")


def test_moe_stages_stay_inside_their_outputs():
    T, E, topk, H, I = 37, 8, 2, 512, 256
    g = torch.Generator().manual_seed(9)
    ids = torch.randint(0, E, (T, topk), generator=g, dtype=torch.int32).to(DEV)
    n = T * topk
    P = int(capi.lib.b2_moe_max_padded(n, E))
    sorted_ids, tile_expert = Guarded(1, P, torch.int32), Guarded(1, P // 128, torch.int32)
    inv_pos, npp = Guarded(1, n, torch.int32), Guarded(1, 1, torch.int32)
    capi.check(capi.lib.b2_moe_align_block128(ids.data_ptr(), n, E, sorted_ids.t.data_ptr(), tile_expert.t.data_ptr(),
                                              inv_pos.t.data_ptr(), npp.t.data_ptr(), P, _st()))
    for gd, name in ((sorted_ids, "sorted_ids"), (tile_expert, "tile_expert"), (inv_pos, "inv_pos"), (npp, "num_post_pad")):
        gd.check("moe_align " + name)
    x = (torch.randn(T, H, generator=g) / 10).bfloat16().to(DEV)
    aq, a_s = Guarded(P, H, torch.int8), Guarded(1, P, torch.float32)
    capi.check(capi.lib.b2_quant_int8_rows_gather(x.data_ptr(), capi.DTYPE_BF16, H, sorted_ids.t.data_ptr(), topk, n,
                                                  aq.t.data_ptr(), H, a_s.t.data_ptr(), P, H, _st()))
    aq.check("gather quant codes"); a_s.check("gather quant scales")
    w = torch.randint(-127, 128, (E, 2 * I, H), generator=g, dtype=torch.int8).to(DEV)
    ws = (torch.rand(E, 2 * I, generator=g) * 1e-2).to(DEV)
    D = Guarded(P, 2 * I, torch.bfloat16)
    capi.check(capi.lib.b2_gemm_s8_grouped(aq.t.data_ptr(), H, w.data_ptr(), H, a_s.t.data_ptr(), ws.data_ptr(),
                                           tile_expert.t.data_ptr(), E, D.t.data_ptr(), 2 * I, capi.DTYPE_BF16, P, 2 * I, H,
                                           _st()))
    D.check("grouped GEMM output")
    tw = torch.rand(T, topk, generator=g).to(DEV)
    y = torch.randn(P, H, device=DEV).bfloat16()
    out = Guarded(T, H, torch.bfloat16, ld=H + 8)
    capi.check(capi.lib.b2_moe_combine(y.data_ptr(), H, inv_pos.t.data_ptr(), tw.data_ptr(), out.t.data_ptr(), H + 8, T,
                                       topk, H, capi.DTYPE_BF16, _st()))
    out.check("moe combine output")
