"""Pins the oracle (oracle/restate.c) before anything trusts it:

  1. against the golden vectors generated from the reference's own compiled CPU code
     (tests/golden/*.npz, made by tests/golden/make_golden.py) -- runs anywhere;
  2. against that compiled reference live (oracle/_ref/sgl_cpu_oracle.so) on the reference's own
     test grid -- only on hosts with AVX512-BF16 + AMX where the .so was built or shipped;
  3. the e4m3 codec against torch's float8_e4m3fn conversions (an independent implementation);
  4. the fp8 restatements against plain fp32 torch math.
"""
from __future__ import annotations

import pytest
import torch

from conftest import bf16_bits, golden_names, load_golden, mk_fp8_block_weight, mk_fp8_weight, mk_int8_problem
from oracle import build_ref, restate


# ------------------------------------------------------------- golden vectors ----

@pytest.mark.parametrize("name", golden_names("int8_"))
def test_int8_restatement_matches_reference_golden(name):
    g = load_golden(name)
    q, s = restate.per_token_quant_int8_cpu(g["A"])
    assert torch.equal(q, g["Aq_u8"]), "u8 codes differ from the reference"
    assert torch.equal(s.view(torch.int32), g["As"].view(torch.int32)), "row scales differ (bitwise)"
    bias = g.get("bias")
    out = restate.int8_scaled_mm_cpu(g["Aq_u8"], g["W"], g["As"], g["Bs"], bias)
    assert torch.equal(bf16_bits(out), bf16_bits(g["out"])), "bf16 output differs from the reference (bitwise)"
    fused = restate.int8_scaled_mm_with_quant(g["A"], g["W"], g["Bs"], bias)
    assert torch.equal(bf16_bits(fused), bf16_bits(g["out"]))


def test_golden_edge_case_covers_eps_and_extremes():
    g = load_golden("int8_edge_M8_N64_K512_bias1")
    # all-zero row -> amax clamps to 1e-7, codes are exactly 128 (= 0 signed)
    assert torch.all(g["Aq_u8"][0] == 128)
    assert g["As"][0].item() == pytest.approx(1e-7 / 127, rel=1e-6)
    # outlier row reaches both ends of the code range
    assert g["Aq_u8"][1].max().item() == 255 and g["Aq_u8"][1].min().item() == 1
    # exact ties round half to EVEN: 0.5*127 = 63.5 -> 64
    assert torch.all(g["Aq_u8"][3, ::2] == 128 + 64)
    assert g["W"].min().item() == -128 and g["W"].max().item() == 127


@pytest.mark.parametrize("name", golden_names("fp8w8a16_"))
def test_fp8_block_restatement_close_to_reference_w8a16(name):
    """fp8_scaled_mm_cpu keeps activations in bf16 (W8A16).  The W8A8 restatement
    additionally rounds every activation to e4m3 (3 mantissa bits: relative rounding error uniform within
    +-2^-4, RMS ~2.6 %), and that error passes undiminished to the output (independent per element), so the
    two agree only to ~2.5 % in Frobenius norm BY CONSTRUCTION; the bound here is 4 %.  The dequantised-weight
    fp32 matmul below (what the reference's own test compares with, tolerance 1e-2) pins the weight side."""
    g = load_golden(name)
    q, sfa = restate.per_token_group_quant_fp8(g["A"], 128, 1e-10, True)
    out = restate.fp8_blockwise_scaled_mm(q, g["W"], sfa, g["sfb"])
    if "bias" in g:
        out = out + g["bias"].view(1, -1)
    ref = g["out"].float()
    err = (out - ref).norm() / ref.norm()
    assert err < 4e-2, f"relative Frobenius error {err:.3g}"
    # and the dequantised-weight matmul in fp32 (what the reference's own test compares with) is closer still
    lut = restate.e4m3_decode_table()
    Wd = lut[g["W"].long()] * g["sfb"].repeat_interleave(128, 0).repeat_interleave(128, 1)
    ref32 = g["A"].float() @ Wd.t() + (g["bias"].view(1, -1) if "bias" in g else 0)
    err32 = (ref32 - ref).abs().max() / ref.abs().max()
    assert err32 < 1e-2


# ------------------------------------------------------- live reference (.so) ----

needs_ref = pytest.mark.skipif(not build_ref.load(), reason="compiled reference CPU path not available on this host")


@needs_ref
@pytest.mark.parametrize("M", [1, 2, 4, 5, 128, 130])
@pytest.mark.parametrize("has_bias", [False, True])
def test_int8_restatement_matches_live_reference(M, has_bias):
    """Reference grid (test/srt/cpu/test_gemm.py: M in {2,128}, N=384, K=544) widened over the
    M<=4 VNNI / M>4 AMX switch (gemm.h:31-34)."""
    ref = torch.ops.sgl_kernel_ref
    x, W, Bs, b = mk_int8_problem(M, 384, 544, seed=1234 + M, bias=has_bias)
    bf = b.float() if has_bias else None
    q_ref, s_ref = ref.per_token_quant_int8_cpu(x)
    q, s = restate.per_token_quant_int8_cpu(x)
    assert torch.equal(q, q_ref) and torch.equal(s.view(torch.int32), s_ref.view(torch.int32))
    o_ref = ref.int8_scaled_mm_with_quant(x, ref.convert_weight_packed(W), Bs, bf, torch.bfloat16, True)
    o = restate.int8_scaled_mm_with_quant(x, W, Bs, bf)
    assert torch.equal(bf16_bits(o), bf16_bits(o_ref))


@needs_ref
def test_int8_restatement_matches_live_reference_c1():
    """BASELINE config C1: M=128 K=4096 N=4096."""
    ref = torch.ops.sgl_kernel_ref
    x, W, Bs, b = mk_int8_problem(128, 4096, 4096)
    o_ref = ref.int8_scaled_mm_with_quant(x, ref.convert_weight_packed(W), Bs, b.float(), torch.bfloat16, True)
    o = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    assert torch.equal(bf16_bits(o), bf16_bits(o_ref))


@needs_ref
def test_strided_activations_match_live_reference():
    """per_token_quant_int8_cpu takes lda = A.stride(0) (gemm_int8.cpp:391)."""
    ref = torch.ops.sgl_kernel_ref
    g = torch.Generator().manual_seed(5)
    big = (torch.randn(16, 1024 + 64, generator=g) / 10).bfloat16()
    x = big[:, :1024]
    q_ref, s_ref = ref.per_token_quant_int8_cpu(x)
    q, s = restate.per_token_quant_int8_cpu(x)
    assert torch.equal(q, q_ref) and torch.equal(s, s_ref)


# ------------------------------------------------------------------ e4m3 codec ----

def test_e4m3_decode_matches_torch():
    lut = restate.e4m3_decode_table()
    t = torch.arange(256, dtype=torch.uint8).view(torch.float8_e4m3fn).float()
    both_nan = lut.isnan() & t.isnan()
    assert torch.equal(lut[~both_nan], t[~both_nan])
    assert both_nan.sum().item() == 2  # 0x7f, 0xff


def test_e4m3_encode_matches_torch_rne_satfinite():
    L = restate.lib()
    g = torch.Generator().manual_seed(3)
    vals = torch.cat([
        torch.randn(4000, generator=g) * 100, torch.randn(2000, generator=g), torch.randn(2000, generator=g) * 0.01,
        torch.tensor([0.0, -0.0, 448.0, -448.0, 447.9, 464.0, 1e9, -1e9, 2.0 ** -9, 2.0 ** -10, 3 * 2.0 ** -10,
                      2.0 ** -6, 0.0175, 1.0625, 1.1875]),
        restate.e4m3_decode_table()[:127], (restate.e4m3_decode_table()[:126] + restate.e4m3_decode_table()[1:127]) / 2,
    ])
    want = vals.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)  # torch: RNE; clamp == satfinite
    got = torch.tensor([L.orc_f32_to_e4m3(float(v)) for v in vals.tolist()], dtype=torch.uint8)
    assert torch.equal(got, want)


# --------------------------------------------------- fp8 restatements vs torch ----

def test_fp8_row_quant_restatement_vs_torch():
    g = torch.Generator().manual_seed(9)
    x = (torch.randn(37, 1024, generator=g) / 10).bfloat16()
    x[5].zero_()
    q, s = restate.per_token_quant_fp8(x)
    amax = x.float().abs().amax(dim=1)
    assert torch.equal(s, amax / 448.0)
    inv = torch.where(s == 0, torch.zeros_like(s), torch.ones_like(s) / s)
    want = (x.float() * inv[:, None]).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
    assert torch.equal(q, want)
    assert torch.all(q[5] == 0) and s[5] == 0


def test_fp8_group_quant_restatement_vs_torch():
    g = torch.Generator().manual_seed(10)
    x = (torch.randn(19, 512, generator=g) / 10).bfloat16()
    x[3, 128:256].zero_()
    for colmajor in (True, False):
        q, s = restate.per_token_group_quant_fp8(x, 128, 1e-10, colmajor)
        assert s.shape == (19, 4)
        assert s.stride() == ((1, 19) if colmajor else (4, 1))
        xg = x.float().view(19, 4, 128)
        amax = xg.abs().amax(dim=2).clamp_min(1e-10)
        ys = amax / 448.0
        assert torch.equal(s.contiguous(), ys)
        want = (xg / ys[:, :, None]).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8).view(19, 512)
        assert torch.equal(q, want)


def test_fp8_gemm_restatements_vs_torch_fp64():
    lut = restate.e4m3_decode_table().double()
    g = torch.Generator().manual_seed(12)
    x = (torch.randn(9, 512, generator=g) / 10).bfloat16()
    Wq, ws = mk_fp8_weight(64, 512)
    q, s = restate.per_token_quant_fp8(x)
    bias = torch.randn(64, generator=g)
    out = restate.fp8_scaled_mm(q, Wq.view(torch.uint8), s, ws, bias)
    ref = (lut[q.long()] @ lut[Wq.view(torch.uint8).long()].t()) * s.double()[:, None] * ws.double().view(1, -1) + bias
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-6)

    Wb, sfb = mk_fp8_block_weight(256, 512)
    qg, sfa = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    outb = restate.fp8_blockwise_scaled_mm(qg, Wb.view(torch.uint8), sfa, sfb)
    A = lut[qg.long()].view(9, 4, 128) * sfa.double()[:, :, None]
    Wd = lut[Wb.view(torch.uint8).long()] * sfb.double().repeat_interleave(128, 0).repeat_interleave(128, 1)
    refb = A.view(9, 512) @ Wd.t()
    torch.testing.assert_close(outb.double(), refb, rtol=1e-5, atol=1e-6)


def test_int8_acc_is_exact_s8_dot():
    g = torch.Generator().manual_seed(4)
    A = torch.randint(0, 256, (7, 96), dtype=torch.int32, generator=g).to(torch.uint8)
    W = torch.randint(-128, 128, (16, 96), dtype=torch.int32, generator=g).to(torch.int8)
    acc = restate.int8_acc(A, W)
    want = ((A.to(torch.int64) - 128) @ W.to(torch.int64).t()).to(torch.int32)
    assert torch.equal(acc, want)
