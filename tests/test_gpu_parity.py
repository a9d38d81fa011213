"""GPU parity tests (-m gpu): the CUDA path vs the oracle on the same seeded inputs, called through
the C ABI (ctypes, exactly what a cgo/JNI/FFI host would bind) and through the torch ops registered
over it.  Bars: integer codes, scales, int32 accumulators and the int8 path's bf16 outputs are
BIT-EXACT; e4m3 codes are bit-exact against the IEEE restatement; fp8 GEMM outputs are held to the
reference's own test tolerance (sgl-kernel/tests/test_fp8_gemm.py:33-35: rtol 0.02, atol 1 on +-448
inputs) and to 1 bf16 ulp of the fp64-accumulated restatement on realistic inputs.
"""
from __future__ import annotations

import pytest
import torch

from conftest import (bf16_bits, golden_names, load_golden, mk_fp8_block_weight, mk_fp8_weight, mk_int8_problem)

pytestmark = pytest.mark.gpu

import sglang_diffusion_b200 as b2  # noqa: E402
from oracle import restate  # noqa: E402  (the checker)

capi = b2.capi
DEV = "cuda:0"


def _stream():
    return torch.cuda.current_stream().cuda_stream


def c_quant_int8(x: torch.Tensor):
    M, K = x.shape
    q = torch.empty(M, K, dtype=torch.int8, device=x.device)
    s = torch.empty(M, dtype=torch.float32, device=x.device)
    dt = capi.DTYPE_BF16 if x.dtype == torch.bfloat16 else capi.DTYPE_F16
    capi.check(capi.lib.b2_quant_int8_rows(x.data_ptr(), dt, x.stride(0), q.data_ptr(), K, s.data_ptr(), M, K, _stream()))
    return q, s


def c_gemm_s8(A, W, As, Bs, bias, out_dtype=torch.bfloat16, cfg=0):
    M, K = A.shape
    N = W.shape[0]
    if out_dtype == torch.int32:
        D = torch.empty(M, N, dtype=torch.int32, device=A.device)
        od = capi.DTYPE_S32
    else:
        D = torch.empty(M, N, dtype=out_dtype, device=A.device)
        od = capi.DTYPE_BF16 if out_dtype == torch.bfloat16 else capi.DTYPE_F16
    bd = capi.DTYPE_F32
    if bias is not None:
        bd = {torch.bfloat16: capi.DTYPE_BF16, torch.float16: capi.DTYPE_F16, torch.float32: capi.DTYPE_F32}[bias.dtype]
    capi.set_gemm_config(cfg)
    try:
        capi.check(capi.lib.b2_gemm_s8(A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0),
                                       None if As is None else As.data_ptr(), None if Bs is None else Bs.data_ptr(),
                                       None if bias is None else bias.data_ptr(), bd, D.data_ptr(), N, od, M, N, K,
                                       _stream()))
    finally:
        capi.set_gemm_config(0)
    return D


def exact_acc(A: torch.Tensor, W: torch.Tensor) -> torch.Tensor:
    """Exact int32 accumulators on the GPU in fp64 (|acc| <= 127*128*K << 2^53); independent of our kernels."""
    return (A.double() @ W.double().t()).to(torch.int32)


# ============================================================ int8 quantizer ====

@pytest.mark.parametrize("M,K", [(1, 32), (1, 512), (3, 544), (5, 4096), (128, 4096), (257, 14336), (7, 16384),
                                 (33, 5120), (2, 13824), (1024, 3072), (4, 32768)])
def test_quant_int8_bit_exact(M, K):
    x, _, _, _ = mk_int8_problem(M, 8, K, seed=M * 131 + K)
    q_ref, s_ref = restate.per_token_quant_int8_cpu(x)
    q, s = c_quant_int8(x.to(DEV))
    assert torch.equal((q.cpu().to(torch.int16) + 128).to(torch.uint8), q_ref)
    assert torch.equal(s.cpu().view(torch.int32), s_ref.view(torch.int32))


@pytest.mark.parametrize("name", golden_names("int8_"))
def test_quant_int8_matches_reference_golden(name):
    g = load_golden(name)
    q, s = c_quant_int8(g["A"].to(DEV))
    assert torch.equal((q.cpu().to(torch.int16) + 128).to(torch.uint8), g["Aq_u8"])
    assert torch.equal(s.cpu().view(torch.int32), g["As"].view(torch.int32))


def test_quant_int8_edge_rows_and_strided_input():
    K = 1024
    x = torch.zeros(6, K + 64, dtype=torch.bfloat16)
    x[1, 5] = 3.0e38          # near bf16 max
    x[2] = 1e-8               # below the 1e-7 eps
    x[3, ::2] = 0.5           # ties: 63.5 -> 64 (half to even)
    x[3, 1::2] = -1.0
    x[4] = torch.linspace(-1, 1, K + 64).bfloat16()
    x[5, :K] = -7.0
    x[:, K:] = 1000.0         # the padding columns must not be read
    xv = x[:, :K]
    q_ref, s_ref = restate.per_token_quant_int8_cpu(xv)
    q, s = c_quant_int8(x.to(DEV)[:, :K])
    assert torch.equal((q.cpu().to(torch.int16) + 128).to(torch.uint8), q_ref)
    assert torch.equal(s.cpu().view(torch.int32), s_ref.view(torch.int32))
    assert torch.all(q[0] == 0) and torch.all(q[3, ::2] == 64) and torch.all(q[5] == -127)


def test_quant_int8_python_wrapper_shapes():
    """Signature of int8_kernel.py:59-89: any leading dims, scales [..., 1]."""
    x = (torch.randn(2, 5, 256) / 10).bfloat16().to(DEV)
    q, s = b2.per_token_quant_int8(x)
    assert q.shape == x.shape and q.dtype == torch.int8 and s.shape == (2, 5, 1) and s.dtype == torch.float32
    q_ref, s_ref = restate.per_token_quant_int8_cpu(x.cpu().view(-1, 256))
    assert torch.equal((q.cpu().view(-1, 256).to(torch.int16) + 128).to(torch.uint8), q_ref)


def test_quant_int8_f16_input_supported_on_gpu():
    """The CPU reference aborts on fp16; the GPU op takes it, with the same arithmetic."""
    x = (torch.randn(9, 512) / 10).half()
    q, s = c_quant_int8(x.to(DEV))
    xf = x.float()
    amax = xf.abs().amax(dim=1).clamp_min(1e-7)
    inv = torch.full_like(amax, 127.0) / amax
    want = torch.round(xf * inv[:, None]).to(torch.int8)  # torch.round = half to even
    assert torch.equal(q.cpu(), want)
    assert torch.equal(s.cpu(), amax / 127.0)


# ====================================================== int8 GEMM accumulators ====

@pytest.mark.parametrize("cfg", [1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("M,N,K", [(128, 256, 256), (300, 1000, 544), (1, 264, 16), (129, 8, 4096), (5, 384, 544)])
def test_gemm_s8_raw_accumulators_exact_every_tile_config(cfg, M, N, K):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + cfg)
    A = torch.randint(-128, 128, (M, K), dtype=torch.int8, generator=g).to(DEV)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g).to(DEV)
    D = c_gemm_s8(A, W, None, None, None, torch.int32, cfg)
    assert torch.equal(D, exact_acc(A, W))


def splitk_cfg(bn_idx: int, split: int, bn: int = 0) -> int:
    """forced id of the decode kernel (gemm_splitk_sm100.cu): 0x80000 | (bn / 16 - 1) | (S - 1) << 4;
    bn = 32 << bn_idx unless given"""
    bn = bn or (32 << bn_idx)
    return 0x80000 | (bn // 16 - 1) | ((split - 1) << 4)


@pytest.mark.parametrize("bn_idx", [0, 1, 2, 3])
@pytest.mark.parametrize("split", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("M,N,K", [(1, 264, 2048), (16, 1000, 4096), (33, 384, 1168), (100, 520, 1024), (128, 256, 8192)])
def test_gemm_s8_splitk_cluster_accumulators_exact(bn_idx, split, M, N, K):
    """Decode kernel: every tile width x cluster size, K tails (1168 = 9.125 stages), N tails, M up to 128.
    int32 partial sums over the cluster are exact, so the result equals the unsplit dot bit for bit."""
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + bn_idx * 11 + split)
    A = torch.randint(-128, 128, (M, K), dtype=torch.int8, generator=g).to(DEV)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g).to(DEV)
    D = c_gemm_s8(A, W, None, None, None, torch.int32, splitk_cfg(bn_idx, split))
    assert capi.last_gemm_config().startswith("splitk_")
    assert torch.equal(D, exact_acc(A, W))


@pytest.mark.parametrize("amc", [1, 2])  # log2 of the activation-multicast group: 2 or 4 adjacent tiles
@pytest.mark.parametrize("bn_idx,split", [(0, 1), (0, 2), (1, 1), (1, 2), (1, 4), (2, 2)])
@pytest.mark.parametrize("M,N,K", [(64, 1024, 1168), (128, 2048, 4096), (32, 1536, 640), (40, 512, 2048), (100, 520, 1024)])
def test_gemm_s8_splitk_activation_multicast_exact(amc, bn_idx, split, M, N, K):
    """Decode kernel with the activation tile of every stage loaded once per group of 2 / 4 adjacent tiles (each CTA
    fetches a slice and multicasts it; stages are released by a multicast commit).  Developer option -- measured
    slower than independent loads (profiles/r02x_amc_sweep.log), the planner never picks it -- kept exact.  Shapes
    whose tile count or row count does not divide fall back to smaller groups inside the launcher."""
    g = torch.Generator().manual_seed(M * 5 + N + K + bn_idx + split + amc)
    A = torch.randint(-128, 128, (M, K), dtype=torch.int8, generator=g).to(DEV)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g).to(DEV)
    D = c_gemm_s8(A, W, None, None, None, torch.int32, splitk_cfg(bn_idx, split) | (amc << 13))
    assert capi.last_gemm_config().startswith("splitk_")
    assert torch.equal(D, exact_acc(A, W))


@pytest.mark.parametrize("bn", [16, 48, 80, 112, 176, 208, 240])
@pytest.mark.parametrize("split", [1, 4])
@pytest.mark.parametrize("M,N,K", [(5, 1000, 2048), (64, 1296, 1168)])
def test_gemm_s8_splitk_odd_tile_widths_exact(bn, split, M, N, K):
    """Tile widths that are not powers of two (the planner picks e.g. 208 rows for N = 28672): 16-column TMEM
    reads, partial rows padded to 32 words, N tails inside a tile."""
    g = torch.Generator().manual_seed(bn * 13 + split + M)
    A = torch.randint(-128, 128, (M, K), dtype=torch.int8, generator=g).to(DEV)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g).to(DEV)
    D = c_gemm_s8(A, W, None, None, None, torch.int32, splitk_cfg(0, split, bn))
    assert f"x{bn}_s" in capi.last_gemm_config()
    assert torch.equal(D, exact_acc(A, W))


@pytest.mark.parametrize("M", [1, 16, 64, 128])
@pytest.mark.parametrize("K,N", [(4096, 6144), (4096, 28672), (14336, 4096)])
def test_decode_shapes_auto_plan_bitwise_vs_persistent_kernel(M, K, N):
    """BASELINE C2 decode shapes: the planned split-K launch equals the persistent kernel (forced config 1) bit for
    bit, bias and scales included."""
    x, W, Bs, bias = mk_int8_problem(M, N, K, seed=M + N)
    x, W, Bs, bias = x.to(DEV), W.to(DEV), Bs.to(DEV), bias.to(DEV)
    q, s = c_quant_int8(x)
    a = c_gemm_s8(q, W, s, Bs, bias)
    assert capi.last_gemm_config().startswith("splitk_") == (M <= 32 or (M <= 64 and N <= 64 * 148 and K >= 2048) or N < 32 * 148)
    if not capi.last_gemm_config().startswith("splitk_"):
        a = c_gemm_s8(q, W, s, Bs, bias, cfg=splitk_cfg(1, 2))
    b = c_gemm_s8(q, W, s, Bs, bias, cfg=1)
    assert capi.last_gemm_config() == "1cta_128x32"
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))


@pytest.mark.parametrize("graph", [False, True])
def test_static_weight_prefetch_chain_is_bit_identical(graph):
    """b2_set_weight_prefetch: quantizer -> decode GEMM launched programmatically (weights requested under the
    quantizer's tail, activations behind griddepcontrol.wait), 24 layers back to back, eager and graph-replayed:
    the outputs equal the plainly-launched chain bit for bit."""
    M, K, N = 16, 4096, 6144
    layers = []
    for i in range(6):
        x, W, Bs, bias = mk_int8_problem(M, N, K, seed=100 + i)
        layers.append((x.to(DEV), W.to(DEV), Bs.to(DEV), bias.to(DEV)))

    def chain():
        outs = []
        for rep in range(4):
            for x, W, Bs, bias in layers:
                q, s = b2.per_token_quant_int8(x)
                outs.append(b2.int8_scaled_mm(q, W.t(), s, Bs.view(-1, 1), torch.bfloat16, bias))
        return outs

    want = chain()
    torch.cuda.synchronize()
    with capi.static_weights():
        assert capi.lib.b2_get_weight_prefetch() == 1
        if graph:
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                chain()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    got = chain()
            torch.cuda.current_stream().wait_stream(st)
            for _ in range(3):
                g.replay()
        else:
            got = chain()
    assert capi.lib.b2_get_weight_prefetch() == 0
    torch.cuda.synchronize()
    for a, b in zip(want, got):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))


@pytest.mark.parametrize("M", [16, 256])
@pytest.mark.parametrize("graph", [False, True])
def test_static_weight_chain_with_data_dependent_layers(graph, M):
    """Each layer's activations are the previous layer's output: the programmatically launched quantizer must wait
    for the GEMM before it (griddepcontrol.wait) and the GEMM for its quantizer; any early read shows up as a
    different bit pattern after 12 dependent layers.  M = 16 runs the decode kernel, M = 256 the persistent one (also
    launched programmatically under the static-weight declaration)."""
    K = 4096
    g = torch.Generator().manual_seed(5)
    x0 = (torch.randn(M, K, generator=g) / 10).bfloat16().to(DEV)
    Ws = [torch.randint(-127, 128, (K, K), dtype=torch.int8, generator=g).to(DEV) for _ in range(4)]
    Bs = [(torch.rand(K, generator=g) * 2e-4).to(DEV) for _ in range(4)]

    def chain():
        x = x0
        for rep in range(3):
            for W, b in zip(Ws, Bs):
                q, s = b2.per_token_quant_int8(x)
                x = b2.int8_scaled_mm(q, W.t(), s, b.view(-1, 1), torch.bfloat16, None)
        return x

    want = chain()
    torch.cuda.synchronize()
    assert torch.isfinite(want.float()).all() and want.float().abs().max() > 0
    with capi.static_weights():
        if graph:
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                chain()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=st):
                    got = chain()
            torch.cuda.current_stream().wait_stream(st)
            for _ in range(3):
                gr.replay()
        else:
            for _ in range(3):
                got = chain()
    torch.cuda.synchronize()
    assert torch.equal(want.view(torch.int16), got.view(torch.int16))


def test_gemm_s8_extreme_codes_do_not_saturate():
    """K = 16384 of (-128)*(-128) = 2.7e8 < 2^31: the accumulator must be exact, not saturated."""
    M, N, K = 130, 136, 16384
    A = torch.full((M, K), -128, dtype=torch.int8, device=DEV)
    W = torch.full((N, K), -128, dtype=torch.int8, device=DEV)
    W[1] = 127
    D = c_gemm_s8(A, W, None, None, None, torch.int32)
    assert torch.all(D[:, 0] == 128 * 128 * K) and torch.all(D[:, 1] == -128 * 127 * K)


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8, 16, 17, 127, 128, 129, 255, 256, 257, 1000])
def test_gemm_s8_accumulators_m_sweep_auto_config(M):
    N, K = 1536, 1024
    g = torch.Generator().manual_seed(M)
    A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g).to(DEV)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g).to(DEV)
    assert torch.equal(c_gemm_s8(A, W, None, None, None, torch.int32), exact_acc(A, W))


def test_gemm_s8_strided_a_and_w():
    """mat_a may have stride(0) != K (int8_gemm_kernel.cu:674 only checks stride(1) == 1)."""
    M, N, K = 70, 200, 512
    g = torch.Generator().manual_seed(3)
    Ab = torch.randint(-127, 128, (M, K + 128), dtype=torch.int8, generator=g).to(DEV)
    Wb = torch.randint(-127, 128, (N, K + 64), dtype=torch.int8, generator=g).to(DEV)
    A, W = Ab[:, :K], Wb[:, :K]
    assert torch.equal(c_gemm_s8(A, W, None, None, None, torch.int32), exact_acc(A, W))


# ============================================ int8 linear: bf16 output bit-exact ====

@pytest.mark.parametrize("name", golden_names("int8_"))
@pytest.mark.parametrize("fused", [False, True])
def test_int8_linear_matches_reference_golden_bitwise(name, fused):
    g = load_golden(name)
    x, W, Bs = g["A"].to(DEV), g["W"].to(DEV), g["Bs"].to(DEV)
    bias = g["bias"].to(DEV) if "bias" in g else None  # fp32, as the CPU op takes it
    if fused:
        out = b2.int8_scaled_mm_with_quant(x, W, Bs, bias, torch.bfloat16, False)
    else:
        q, s = c_quant_int8(x)
        out = c_gemm_s8(q, W, s, Bs, bias)
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(g["out"]))


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 127, 128, 130])
@pytest.mark.parametrize("has_bias", [False, True])
def test_int8_linear_reference_test_grid_bitwise(M, has_bias):
    """test/srt/cpu/test_gemm.py grid: N=384, K=544 (K % 128 != 0 exercises the TMA K tail)."""
    x, W, Bs, b = mk_int8_problem(M, 384, 544, seed=1234 + M, bias=has_bias)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float() if has_bias else None)
    out = b2.int8_scaled_mm_with_quant(x.to(DEV), W.to(DEV), Bs.to(DEV), b.to(DEV) if has_bias else None,
                                       torch.bfloat16, False)
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref))


@pytest.mark.parametrize("cfg", [1, 2, 3, 4, 5, 6])
def test_int8_c1_bitwise_every_tile_config(cfg):
    """BASELINE config C1: M=128 K=4096 N=4096, bf16 out, bias."""
    x, W, Bs, b = mk_int8_problem(128, 4096, 4096)
    ref, q_ref, s_ref, acc_ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float(), return_parts=True)
    q, s = c_quant_int8(x.to(DEV))
    Wd = W.to(DEV)
    acc = c_gemm_s8(q, Wd, None, None, None, torch.int32, cfg)
    assert torch.equal(acc.cpu(), acc_ref)
    out = c_gemm_s8(q, Wd, s, Bs.to(DEV), b.to(DEV), torch.bfloat16, cfg)
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref))


def test_int8_scaled_mm_op_reference_call_convention():
    """The call W8A8Int8LinearMethod.apply makes (w8a8_int8.py:216-231): mat_b = weight.t(), scales [M,1]/[N,1]."""
    x, W, Bs, b = mk_int8_problem(33, 160, 1024, seed=5)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    layer = torch.nn.Module()
    m = b2.W8A8Int8LinearMethod()
    m.create_weights(layer, 1024, [160], 1024, 160, torch.bfloat16, device=DEV)
    layer.weight.data.copy_(W)
    layer.weight_scale.data.copy_(Bs.view(-1, 1))
    m.process_weights_after_loading(layer)
    out = m.apply(layer, x.to(DEV).view(3, 11, 1024), b.to(DEV))
    assert out.shape == (3, 11, 160)
    assert torch.equal(bf16_bits(out.cpu().view(33, 160)), bf16_bits(ref))
    out_f = b2.W8A8Int8LinearMethod(fused=True).apply(layer, x.to(DEV), b.to(DEV))
    assert torch.equal(bf16_bits(out_f.cpu()), bf16_bits(ref))


def test_int8_f16_output_epilogue():
    x, W, Bs, b = mk_int8_problem(17, 64, 256, seed=8)
    xh = x.half()
    q, s = c_quant_int8(xh.to(DEV))
    out = c_gemm_s8(q, W.to(DEV), s, Bs.to(DEV), b.half().to(DEV), torch.float16)
    acc = exact_acc(q, W.to(DEV)).float()
    want = torch.addcmul(b.half().float().to(DEV).double(), (acc * s[:, None]).double(), Bs.to(DEV).double()[None, :])
    # fma(t, Bs, bias) rounded once to fp32 == fp64 addcmul rounded to fp32 except for double rounding (rare): allow 1 ulp of f16
    torch.testing.assert_close(out.float(), want.float().half().float(), rtol=2e-3, atol=1e-3)


# =========================================== full BASELINE sizes: property tests ====

@pytest.mark.parametrize("K,N", [(4096, 6144), (4096, 28672), (14336, 4096)])
def test_llama8b_shapes_m8192_checksum_and_sampled_rows(K, N):
    """C2 at M=8192.  Size-independent properties: (1) column checksum -- sum_n acc[m,n] == A[m,:] . (sum_n W[n,:])
    exactly, in int64; (2) 48 sampled rows bit-exact against the CPU oracle end to end (codes, scales, bf16)."""
    M = 8192
    g = torch.Generator(device=DEV).manual_seed(K + N)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    x[:, ::512] *= 30  # outlier channels
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g, device=DEV)
    Bs = torch.rand(N, generator=g, device=DEV) * 1e-2
    bias = torch.randn(N, generator=g, device=DEV).bfloat16()
    q, s = c_quant_int8(x)
    acc = c_gemm_s8(q, W, None, None, None, torch.int32)
    wsum = W.double().sum(dim=0)                       # exact: |.| <= 128*28672
    want = (q.double() @ wsum).to(torch.int64)         # exact: < 2^53
    got = acc.to(torch.int64).sum(dim=1)
    assert torch.equal(got, want)
    out = b2.int8_scaled_mm_with_quant(x, W, Bs, bias, torch.bfloat16, False)
    rows = torch.randperm(M, generator=torch.Generator().manual_seed(1))[:48].sort().values
    ref, q_ref, s_ref, acc_ref = restate.int8_scaled_mm_with_quant(x[rows.to(DEV)].cpu(), W.cpu(), Bs.cpu(),
                                                                   bias.float().cpu(), return_parts=True)
    assert torch.equal((q[rows.to(DEV)].cpu().to(torch.int16) + 128).to(torch.uint8), q_ref)
    assert torch.equal(acc[rows.to(DEV)].cpu(), acc_ref)
    assert torch.equal(bf16_bits(out[rows.to(DEV)].cpu()), bf16_bits(ref))


@pytest.mark.parametrize("K,N", [(5120, 5120), (5120, 13824), (13824, 5120), (3072, 9216), (3072, 12288), (12288, 3072),
                                 (15360, 3072)])
def test_dit_shapes_m32768_checksum_and_sampled_rows(K, N):
    """C4 at the full diffusion token batch M = 32768 (Wan2.1-14B 5120/13824, FLUX.1 3072/9216/12288/15360): the column
    checksum of the raw accumulators is exact and 32 sampled rows equal the CPU oracle bit for bit, bias included."""
    M = 32768
    g = torch.Generator(device=DEV).manual_seed(K * 3 + N)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g, device=DEV)
    Bs = torch.rand(N, generator=g, device=DEV) * 1e-2
    bias = torch.randn(N, generator=g, device=DEV).bfloat16()
    q, s = c_quant_int8(x)
    acc = c_gemm_s8(q, W, None, None, None, torch.int32)
    want = (q.double() @ W.double().sum(dim=0)).to(torch.int64)
    assert torch.equal(acc.to(torch.int64).sum(dim=1), want)
    del acc
    out = b2.int8_scaled_mm(q, W.t(), s.view(-1, 1), Bs.view(-1, 1), torch.bfloat16, bias)
    rows = torch.randperm(M, generator=torch.Generator().manual_seed(2))[:32].sort().values
    ref = restate.int8_scaled_mm_with_quant(x[rows.to(DEV)].cpu(), W.cpu(), Bs.cpu(), bias.float().cpu())
    assert torch.equal(bf16_bits(out[rows.to(DEV)].cpu()), bf16_bits(ref))


@pytest.mark.parametrize("K,N", [(8192, 1280), (8192, 7168), (1024, 8192), (3584, 8192)])
@pytest.mark.parametrize("M", [64, 8192])
def test_llama70b_tp8_fp8_shards_sampled_rows(K, N, M):
    """C5: the Llama-3-70B TP = 8 shard shapes at decode (M = 64, cluster split-K kernel) and prefill (M = 8192,
    2-CTA kernel): fp8 codes and scales bit-exact against the oracle, 32 sampled output rows within the reference's
    tolerance of the fp64 restatement."""
    g = torch.Generator(device=DEV).manual_seed(K + N + M)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    Wq, ws = mk_fp8_weight(N, K)
    xq, xs = b2.sglang_per_token_quant_fp8(x)
    out = b2.fp8_scaled_mm(xq, Wq.to(DEV).t(), xs, ws.to(DEV), torch.bfloat16, None)
    rows = torch.randperm(M, generator=torch.Generator().manual_seed(3))[:32].sort().values
    q_ref, s_ref = restate.per_token_quant_fp8(x[rows.to(DEV)].cpu())
    assert torch.equal(xq[rows.to(DEV)].cpu().view(torch.uint8), q_ref)
    assert torch.equal(xs[rows.to(DEV)].cpu().view(-1).view(torch.int32), s_ref.view(-1).view(torch.int32))
    ref = restate.fp8_scaled_mm(q_ref, Wq.view(torch.uint8), s_ref, ws, None)
    got = out[rows.to(DEV)].float().cpu()
    torch.testing.assert_close(got, ref, rtol=1e-2, atol=1e-2 * ref.abs().max().item())


def test_m_sharded_rows_equal_unsharded_bitwise():
    """C4 (DiT, Wan2.1 hidden 5120): splitting M over ranks changes nothing -- rows are independent."""
    M, K, N = 4096, 5120, 5120
    g = torch.Generator(device=DEV).manual_seed(4)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=DEV)
    Bs = torch.rand(N, generator=g, device=DEV) * 1e-2
    full = b2.int8_scaled_mm_with_quant(x, W, Bs, None, torch.bfloat16, False)
    for world in (2, 8):
        parts = []
        for r in range(world):
            a, b = b2.shard_rows(M, r, world)
            parts.append(b2.int8_scaled_mm_with_quant(x[a:b], W, Bs, None, torch.bfloat16, False))
        assert torch.equal(bf16_bits(torch.cat(parts)), bf16_bits(full))


def test_gemm_s8_linearity_in_weights():
    """acc(A, W1 + W2) == acc(A, W1) + acc(A, W2) exactly (int32), at a DiT shape."""
    M, K, N = 2048, 3072, 9216
    g = torch.Generator(device=DEV).manual_seed(6)
    A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g, device=DEV)
    W1 = torch.randint(-60, 61, (N, K), dtype=torch.int8, generator=g, device=DEV)
    W2 = torch.randint(-60, 61, (N, K), dtype=torch.int8, generator=g, device=DEV)
    a12 = c_gemm_s8(A, W1 + W2, None, None, None, torch.int32)
    assert torch.equal(a12, c_gemm_s8(A, W1, None, None, None, torch.int32) + c_gemm_s8(A, W2, None, None, None, torch.int32))


# ================================================================ fp8 quantizers ====

@pytest.mark.parametrize("M", [128, 256, 512])
@pytest.mark.parametrize("K", [512, 1076, 1368, 2048, 4096])
def test_per_token_quant_fp8_reference_grid(M, K):
    """Grid of sgl-kernel/tests/test_per_token_quant_fp8.py (1076/1368 are the vec-4 shapes); bit-exact vs the
    IEEE restatement, which is tighter than the reference's own 1e-3 check."""
    g = torch.Generator().manual_seed(M + K)
    x = torch.randn(M, K, generator=g).bfloat16()
    q_ref, s_ref = restate.per_token_quant_fp8(x)
    q, s = b2.sglang_per_token_quant_fp8(x.to(DEV))
    assert s.shape == (M, 1)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref)
    assert torch.equal(s.cpu().view(-1).view(torch.int32), s_ref.view(torch.int32))


def test_per_token_quant_fp8_zero_row_and_saturation():
    x = torch.zeros(4, 256, dtype=torch.bfloat16)
    x[1, 3] = 5.0
    x[2] = -2.0
    x[3, 0] = 3.0e38
    q, s = b2.sglang_per_token_quant_fp8(x.to(DEV))
    q8 = q.cpu().view(torch.uint8)
    assert s[0].item() == 0 and torch.all(q8[0] == 0)            # per_token_quant_fp8.cu: s == 0 -> inv = 0
    assert q8[1, 3].item() == 0x7E and q8[2, 0].item() == 0xFE   # +-448 codes
    q_ref, s_ref = restate.per_token_quant_fp8(x)
    assert torch.equal(q8, q_ref) and torch.equal(s.cpu().view(-1), s_ref)


@pytest.mark.parametrize("M,K", [(1, 128), (1, 512), (3, 7168), (127, 7168), (512, 2048), (4096, 1536)])
@pytest.mark.parametrize("colmajor", [True, False])
def test_group_quant_fp8_128_bit_exact(M, K, colmajor):
    g = torch.Generator().manual_seed(M * 3 + K)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    if M > 2:
        x[2, :128].zero_()   # eps path
    q_ref, s_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, colmajor)
    q, s = b2.per_token_group_quant_fp8(x.to(DEV), 128, column_major_scales=colmajor)
    assert s.shape == (M, K // 128)
    if colmajor and M > 1 and K > 128:
        assert s.stride() == (1, M)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref)
    assert torch.equal(s.cpu().contiguous().view(torch.int32), s_ref.contiguous().view(torch.int32))


def test_group_quant_fp8_shared_reciprocal_division_is_exact_for_every_significand_pair():
    """The group quantizer replaces x / ys by Markstein's reciprocal sequence (csrc/quant.cu).  Every bf16
    significand of x (128) x sign x exponent gap (0..15) against every bf16 significand of the group maximum (128)
    at four magnitudes: codes and scales must equal the oracle's true division bit for bit."""
    sig = (128 + torch.arange(128, dtype=torch.float64)) / 128.0  # 1.0000000b .. 1.1111111b
    xs = []
    for gap in range(16):
        for sign in (1.0, -1.0):
            xs.append(sign * sig * 2.0 ** (-gap))
    xs = torch.cat(xs)  # 4096 values relative to 2^ea
    per_group = 127
    ngroups = (xs.numel() + per_group - 1) // per_group
    pad = ngroups * per_group - xs.numel()
    xs = torch.cat([xs, xs[:pad]]).view(ngroups, per_group)
    rows = []
    for ea in (-20, 0, 7, 60):
        for a in sig.tolist():
            blk = torch.empty(ngroups, 128, dtype=torch.float64)
            blk[:, 0] = a * 2.0 ** ea
            blk[:, 1:] = xs * 2.0 ** ea
            rows.append(blk)
    x = torch.cat(rows).to(torch.bfloat16)  # every value is exactly representable in bf16
    assert x.shape[0] == 4 * 128 * ngroups
    q_ref, s_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    q, s = b2.per_token_group_quant_fp8(x.to(DEV), 128, column_major_scales=True)
    assert torch.equal(s.cpu().contiguous().view(torch.int32), s_ref.contiguous().view(torch.int32))
    assert torch.equal(q.cpu().view(torch.uint8), q_ref)
    # independent of the oracle: torch's own fp32 division and e4m3 cast on the device
    xf = x.to(DEV).float()
    want = (xf / s.reshape(-1, 1)).clamp(-448, 448).to(torch.float8_e4m3fn)
    assert torch.equal(q.view(torch.uint8), want.view(torch.uint8))


# ===================================================================== fp8 GEMMs ====

def _ulp_bf16(t: torch.Tensor) -> torch.Tensor:
    return torch.maximum(t.abs(), torch.full_like(t, 2.0 ** -126)).log2().floor().exp2() * 2.0 ** -7


@pytest.mark.parametrize("M,N,K", [(1, 512, 512), (5, 128, 512), (64, 1280, 8192), (64, 8192, 1024),
                                   (300, 7168, 1024), (1024, 8192, 3584)])
@pytest.mark.parametrize("has_bias", [False, True])
def test_fp8_scaled_mm_vs_restatement(M, N, K, has_bias):
    """C5 shard shapes.  fp32 tensor-core accumulation vs exact-product fp64 accumulation: within
    1e-2 normalised, and >= 99 % of outputs equal to the bf16 rounding of the exact result (<= 1 ulp all)."""
    g = torch.Generator().manual_seed(M + N + K)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    Wq, ws = mk_fp8_weight(N, K)
    bias = torch.randn(N, generator=g).bfloat16() if has_bias else None
    q_ref, s_ref = restate.per_token_quant_fp8(x)
    ref = restate.fp8_scaled_mm(q_ref, Wq.view(torch.uint8), s_ref, ws, bias.float() if has_bias else None)
    out = b2.apply_fp8_linear(x.to(DEV), Wq.to(DEV).t(), ws.to(DEV), bias=bias.to(DEV) if has_bias else None)
    o = out.float().cpu()
    assert ((o - ref).abs().max() / ref.abs().max()).item() < 1e-2
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 1e-6 * ref.abs().max())
    assert (o == ref.bfloat16().float()).float().mean().item() > 0.99


@pytest.mark.parametrize("bn_idx,split", [(0, 1), (0, 8), (1, 3), (2, 4), (3, 2), (3, 7)])
@pytest.mark.parametrize("M", [1, 64, 128])
def test_fp8_splitk_cluster_vs_restatement(bn_idx, split, M):
    """fp8 decode kernel: fp32 partials summed in rank order; within the reference tolerance and deterministic."""
    N, K = 1288, 3584
    g = torch.Generator(device=DEV).manual_seed(bn_idx * 8 + split + M)
    a = (torch.randn(M, K, generator=g, device=DEV)).clamp(-448, 448).to(torch.float8_e4m3fn)
    b = (torch.randn(N, K, generator=g, device=DEV)).clamp(-448, 448).to(torch.float8_e4m3fn)
    sa = torch.rand(M, generator=g, device=DEV) * 1e-2 + 1e-3
    sb = torch.rand(N, generator=g, device=DEV) * 1e-2 + 1e-3
    bias = torch.randn(N, generator=g, device=DEV).bfloat16()
    outs = []
    for _ in range(2):
        capi.set_gemm_config(splitk_cfg(bn_idx, split))
        try:
            outs.append(b2.fp8_scaled_mm(a, b.t(), sa.view(-1, 1), sb.view(-1, 1), torch.bfloat16, bias))
            assert capi.last_gemm_config().startswith("splitk_")
        finally:
            capi.set_gemm_config(0)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))
    o = (a.double() @ b.double().t()) * sa.double()[:, None] * sb.double()[None, :] + bias.double()
    torch.testing.assert_close(outs[0].double(), o, rtol=1e-2, atol=1e-2 * o.abs().max().item())


@pytest.mark.parametrize("cfg", [1, 2, 3, 4, 5, 6])
def test_fp8_scaled_mm_reference_test_distribution(cfg):
    """sgl-kernel/tests/test_fp8_gemm.py: operands uniform in +-448, scales randn*1e-3... rtol 0.02, atol 1."""
    M, N, K = 257, 1024, 2048
    g = torch.Generator(device=DEV).manual_seed(cfg)
    a = ((torch.rand(M, K, generator=g, device=DEV) * 2 - 1) * 448).to(torch.float8_e4m3fn)
    b = ((torch.rand(N, K, generator=g, device=DEV) * 2 - 1) * 448).to(torch.float8_e4m3fn)
    sa = torch.randn(M, generator=g, device=DEV) * 1e-3
    sb = torch.randn(N, generator=g, device=DEV) * 1e-3
    bias = torch.randn(N, generator=g, device=DEV).bfloat16()
    capi.set_gemm_config(cfg)
    try:
        out = b2.fp8_scaled_mm(a, b.t(), sa.view(-1, 1), sb.view(-1, 1), torch.bfloat16, bias)
    finally:
        capi.set_gemm_config(0)
    o = (a.double() @ b.double().t()) * sa.double()[:, None] * sb.double()[None, :] + bias.double()
    torch.testing.assert_close(out.double(), o, rtol=0.02, atol=1)
    torch.testing.assert_close(out.double(), o, rtol=1e-2, atol=1e-2 * o.abs().max().item())


@pytest.mark.parametrize("M", [1, 3, 5, 127, 128, 512, 1024])
@pytest.mark.parametrize("N,K", [(512, 512), (2048, 7168), (4096, 1536)])
def test_fp8_blockwise_scaled_mm(M, N, K):
    """Grid after sgl-kernel/tests/test_fp8_blockwise_gemm.py (M list incl. non-multiples of 4: no padding copies
    here); scales in the reference's layouts -- SFA M-major, SFB K-major (:76-77)."""
    g = torch.Generator().manual_seed(M + N)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    Wq, sfb = mk_fp8_block_weight(N, K)
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.view(torch.uint8), sfa_ref, sfb)
    out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x.to(DEV), Wq.to(DEV), [128, 128], sfb.to(DEV))
    o = out.float().cpu()
    assert ((o - ref).abs().max() / ref.abs().max()).item() < 1e-2
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 2e-6 * ref.abs().max())


@pytest.mark.parametrize("name", golden_names("fp8w8a16_"))
def test_fp8_blockwise_vs_reference_cpu_golden(name):
    """Against the reference's own fp8_scaled_mm_cpu output (W8A16): agreement is bounded by the e4m3
    rounding of the activations (~2.5 % Frobenius, see tests/test_oracle.py); bound 4 %."""
    g = load_golden(name)
    out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(
        g["A"].to(DEV), g["W"].view(torch.float8_e4m3fn).to(DEV), [128, 128], g["sfb"].to(DEV),
        bias=g["bias"].bfloat16().to(DEV) if "bias" in g else None)
    ref = g["out"].float()
    assert ((out.float().cpu() - ref).norm() / ref.norm()).item() < 4e-2


def _mk_fp8_block_weight_dev(N, K, seed=11):
    """mk_fp8_block_weight on the device (264 M-element C3 weights in milliseconds); returns device e4m3 [N,K]
    and fp32 block scales [N/128, K/128]."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    w = torch.randn(N, K, generator=g, device=DEV) * 0.05
    wb = w.view(N // 128, 128, K // 128, 128)
    sfb = (wb.abs().amax(dim=(1, 3)) / 448.0).float()
    Wq = (wb / sfb[:, None, :, None]).clamp(-448, 448).reshape(N, K).to(torch.float8_e4m3fn)
    return Wq, sfb


C3_SHAPES = [(24576, 1536), (32768, 512), (7168, 16384), (36864, 7168), (7168, 18432), (4096, 7168), (7168, 2048)]


@pytest.mark.parametrize("N,K", C3_SHAPES)
def test_fp8_blockwise_c3_shape_m4096_sampled(N, K):
    """C3 (DeepSeek-V3, M=4096), every (N,K) of sgl-kernel/benchmark/bench_fp8_blockwise_gemm.py:32-53: sampled rows
    (first / last row of the batch, rows across M tiles) against the restatement, every column."""
    M = 4096
    g = torch.Generator(device=DEV).manual_seed(2)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    Wq, sfb = _mk_fp8_block_weight_dev(N, K)
    out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x, Wq, [128, 128], sfb)
    rows = torch.tensor([0, 1, 127, 128, 255, 1337, 2048, 3000, 3967, 4095])
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x[rows.to(DEV)].cpu().contiguous(), 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.cpu().view(torch.uint8), sfa_ref, sfb.cpu())
    o = out[rows.to(DEV)].float().cpu()
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 2e-6 * ref.abs().max())
    # size-independent property: every row of the full output is finite and the column checksum of the sampled rows'
    # neighbours is reproducible (the kernel is deterministic: a second call returns the same bits)
    out2 = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x, Wq, [128, 128], sfb)
    assert torch.equal(out.view(torch.int16), out2.view(torch.int16))
    assert bool(torch.isfinite(out).all())


@pytest.mark.parametrize("cfg,name", [(1, "1cta_128x128"), (7, "1cta_128x64"), (8, "1cta_128x32"), (2, "2cta_256x128")])
@pytest.mark.parametrize("M,N,K", [(1, 4096, 7168), (16, 4096, 7168), (128, 7168, 2048), (5, 1536, 640), (300, 512, 1152)])
def test_fp8_blockwise_every_tile_width(cfg, name, M, N, K):
    """128 / 64 / 32-wide 1-CTA tiles (the narrow ones are the decode-shape schedule: a CTA on every SM, each inside one
    128-column scale block, stores straight from registers) and the 2-CTA tile: all within 1 ulp of the restatement,
    and the narrow tiles bit-identical to the 128-wide one (same per-element arithmetic, K blocks in order)."""
    g = torch.Generator().manual_seed(M + N + K)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    Wq, sfb = mk_fp8_block_weight(N, K)
    bias = torch.randn(N, generator=g).bfloat16()
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.view(torch.uint8), sfa_ref, sfb) + bias.float()
    outs = {}
    for c in (cfg, 1):
        capi.set_gemm_config(c)
        try:
            outs[c] = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x.to(DEV), Wq.to(DEV), [128, 128], sfb.to(DEV),
                                                                     bias=bias.to(DEV))
            if c == cfg:
                assert capi.last_gemm_config().startswith(name), capi.last_gemm_config()
        finally:
            capi.set_gemm_config(0)
    o = outs[cfg].float().cpu()
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 2e-6 * ref.abs().max())
    assert torch.equal(outs[cfg].view(torch.int16), outs[1].view(torch.int16))


@pytest.mark.parametrize("cfg", [9, 25, 26, 41, 42, 43])
@pytest.mark.parametrize("M,N,K", [(1, 4096, 7168), (16, 4096, 7168), (128, 1536, 2048), (37, 512, 1664)])
def test_fp8_blockwise_k_split_cluster(cfg, M, N, K):
    """K split over a {2,4,8}-CTA cluster for the decode tiles (partial sums pushed to rank 0 over DSMEM, added in rank
    order): within 1 ulp of the restatement, deterministic.  cfg = 9 + 16*b + s: width 128 >> b, split 2 << s."""
    b, sidx = (cfg - 9) // 16, (cfg - 9) % 16
    g = torch.Generator().manual_seed(M + N + K + cfg)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    Wq, sfb = mk_fp8_block_weight(N, K)
    bias = torch.randn(N, generator=g).bfloat16()
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.view(torch.uint8), sfa_ref, sfb) + bias.float()
    outs = []
    for _ in range(2):
        capi.set_gemm_config(cfg)
        try:
            outs.append(b2.cutlass_w8a8_block_fp8_linear_with_fallback(x.to(DEV), Wq.to(DEV), [128, 128], sfb.to(DEV),
                                                                       bias=bias.to(DEV)))
            assert capi.last_gemm_config() == f"1cta_128x{128 >> b}_block128_ks{2 << sidx}", capi.last_gemm_config()
        finally:
            capi.set_gemm_config(0)
    assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16))
    o = outs[0].float().cpu()
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 4e-6 * ref.abs().max())


@pytest.mark.parametrize("K", [128, 384, 640, 1664, 2176])
@pytest.mark.parametrize("M,N", [(1, 128), (5, 512), (127, 256), (300, 1536), (1024, 2048)])
def test_fp8_blockwise_odd_k_block_counts(M, N, K):
    """K/128 odd (1, 3, 5, 13, 17 K-blocks): the tail of the two-blocks-per-step promotion loop and a scale chunk
    that is not full (17 = one chunk of 16 + 1)."""
    g = torch.Generator().manual_seed(M + N + K)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    Wq, sfb = mk_fp8_block_weight(N, K)
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.view(torch.uint8), sfa_ref, sfb)
    out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x.to(DEV), Wq.to(DEV), [128, 128], sfb.to(DEV))
    o = out.float().cpu()
    assert ((o - ref).abs().max() / ref.abs().max()).item() < 1e-2
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 2e-6 * ref.abs().max())


@pytest.mark.parametrize("M", [1, 3, 5, 127, 128])
@pytest.mark.parametrize("N,K", [(7168, 16384), (4096, 7168), (7168, 18432)])
def test_fp8_blockwise_small_m_long_k(M, N, K):
    """The reference test's M list (sgl-kernel/tests/test_fp8_blockwise_gemm.py) at the long-K C3 shapes: 128 / 144
    K-blocks wrap the TMEM slot ring 32+ times and span 8-9 scale chunks."""
    g = torch.Generator(device=DEV).manual_seed(M + K)
    x = (torch.randn(M, K, generator=g, device=DEV) / 10).bfloat16()
    Wq, sfb = _mk_fp8_block_weight_dev(N, K)
    out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x, Wq, [128, 128], sfb)
    q_ref, sfa_ref = restate.per_token_group_quant_fp8(x.cpu(), 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.cpu().view(torch.uint8), sfa_ref, sfb.cpu())
    o = out.float().cpu()
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 2e-6 * ref.abs().max())


# ================================================== boundary behaviour on the GPU ====

def test_argument_errors_follow_reference_checks():
    a = torch.zeros(4, 64, dtype=torch.int8, device=DEV)
    w = torch.zeros(32, 64, dtype=torch.int8, device=DEV)
    sa, sb = torch.ones(4, device=DEV), torch.ones(32, device=DEV)
    with pytest.raises(RuntimeError, match="column major"):
        b2.int8_scaled_mm(a, w.t().contiguous(), sa, sb, torch.bfloat16)
    with pytest.raises(RuntimeError, match="Half or BFloat16"):
        b2.int8_scaled_mm(a, w.t(), sa, sb, torch.float32)
    with pytest.raises(RuntimeError, match="scales_a"):
        b2.int8_scaled_mm(a, w.t(), sa[:3], sb, torch.bfloat16)
    with pytest.raises(RuntimeError, match="bias dtype"):
        b2.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, torch.zeros(32, device=DEV))
    with pytest.raises(RuntimeError, match="multiple of 16"):
        b2.int8_scaled_mm(a[:, :40], w[:, :40].t(), sa, sb, torch.bfloat16)
    with pytest.raises(RuntimeError, match="Int8"):
        b2.int8_scaled_mm(a.to(torch.uint8), w.t(), sa, sb, torch.bfloat16)
    # C ABI: status code + message, no exception type
    st = capi.lib.b2_gemm_s8(a.data_ptr(), 64, w.data_ptr(), 64, sa.data_ptr(), sb.data_ptr(), None, 0, a.data_ptr(),
                             30, capi.DTYPE_BF16, 4, 30, 64, _stream())
    assert st == capi.ERR_ARG and "multiple of 8" in capi.last_error()


def test_empty_batch_is_a_noop():
    w = torch.zeros(32, 64, dtype=torch.int8, device=DEV)
    x = torch.zeros(0, 64, dtype=torch.bfloat16, device=DEV)
    capi.reset_launch_count()
    out = b2.int8_scaled_mm_with_quant(x, w, torch.ones(32, device=DEV), None, torch.bfloat16, False)
    assert out.shape == (0, 32) and capi.launch_count() == 0


def test_ops_are_cuda_graph_capturable_and_use_the_current_stream():
    x, W, Bs, b = mk_int8_problem(64, 256, 512, seed=21)
    xd, Wd, Bsd, bd = x.to(DEV), W.to(DEV), Bs.to(DEV), b.to(DEV)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        b2.int8_scaled_mm_with_quant(xd, Wd, Bsd, bd, torch.bfloat16, False)  # warm-up (func attributes)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out = b2.int8_scaled_mm_with_quant(xd, Wd, Bsd, bd, torch.bfloat16, False)
    torch.cuda.current_stream().wait_stream(side)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref))
    xd.copy_((x * 2).to(DEV))
    graph.replay()
    torch.cuda.synchronize()
    ref2 = restate.int8_scaled_mm_with_quant((x * 2), W, Bs, b.float())
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref2))


def test_launch_counter_counts_our_kernels():
    x, W, Bs, b = mk_int8_problem(64, 64, 128, seed=2)
    capi.reset_launch_count()
    b2.int8_scaled_mm_with_quant(x.to(DEV), W.to(DEV), Bs.to(DEV), None, torch.bfloat16, False)
    assert capi.launch_count() == 2  # quantizer + GEMM
    capi.reset_launch_count()
    with capi.fused_decode():
        b2.int8_scaled_mm_with_quant(x[:8].to(DEV), W.to(DEV), Bs.to(DEV), None, torch.bfloat16, False)
    assert capi.launch_count() == 1  # M <= 32, opt-in: the decode GEMM quantizes its activations itself
    assert capi.last_gemm_config().endswith("_fusedq")


@pytest.mark.parametrize("M", [1, 2, 5, 8, 16, 17, 31, 32])
@pytest.mark.parametrize("K,N", [(4096, 6144), (4096, 28672), (14336, 4096), (544, 384), (16, 8), (2048, 1000)])
@pytest.mark.parametrize("has_bias", [False, True])
def test_fused_quant_decode_linear_bitwise_vs_oracle(M, K, N, has_bias):
    """M <= 32: ONE kernel (cluster split-K GEMM whose epilogue warps first quantize the CTA's K slice of the bf16
    activations into the resident A tiles, row maxima exchanged over DSMEM).  Output == oracle bit for bit == the
    two-launch form (B2_FUSED_DECODE=0 equivalent: separate quantizer + GEMM)."""
    x, W, Bs, b = mk_int8_problem(M, N, K, seed=M + K + N, bias=has_bias)
    x[0, ::7] *= 30  # outliers: the row maximum sits in ONE K slice, the others must learn it through the cluster
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float() if has_bias else None)
    bd = b.to(DEV) if has_bias else None
    with capi.static_weights(), capi.fused_decode():
        out = b2.int8_scaled_mm_with_quant(x.to(DEV), W.to(DEV), Bs.to(DEV), bd, torch.bfloat16, False)
    assert capi.last_gemm_config().endswith("_fusedq"), capi.last_gemm_config()
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref))
    q, s = b2.per_token_quant_int8(x.to(DEV))
    two = b2.int8_scaled_mm(q, W.to(DEV).t(), s, Bs.to(DEV).view(-1, 1), torch.bfloat16, bd)
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(two.cpu()))


@pytest.mark.parametrize("bn,split", [(32, 1), (64, 2), (128, 4), (128, 8), (256, 3), (48, 5)])
def test_fused_quant_decode_forced_plans_and_strided_input(bn, split):
    M, N, K = 16, 1000, 4096
    x, W, Bs, b = mk_int8_problem(M, N, K, seed=bn + split)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    xs = torch.zeros(M, K + 64, dtype=torch.bfloat16, device=DEV)  # row stride != K
    xs[:, :K] = x.to(DEV)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    Wd, Bsd, bd = W.to(DEV), Bs.to(DEV), b.to(DEV)
    capi.set_gemm_config(splitk_cfg(0, split, bn))
    capi.lib.b2_set_fused_decode(1)
    try:
        st = capi.lib.b2_linear_w8a8_int8(xs.data_ptr(), capi.DTYPE_BF16, K + 64, Wd.data_ptr(), K, Bsd.data_ptr(),
                                          bd.data_ptr(), capi.DTYPE_BF16, out.data_ptr(), N, M, N, K, None, 0, _stream())
    finally:
        capi.set_gemm_config(0)
        capi.lib.b2_set_fused_decode(0)
    capi.check(st)
    assert capi.last_gemm_config().endswith("_fusedq")
    assert torch.equal(bf16_bits(out.cpu()), bf16_bits(ref))


@pytest.mark.parametrize("M", [1, 16, 32])
@pytest.mark.parametrize("K,N", [(8192, 1280), (1024, 8192), (3584, 8192)])
def test_fused_quant_decode_fp8_linear(M, K, N):
    """fp8 twin (b2_linear_w8a8_fp8): e4m3 codes are bit-identical to sgl_per_token_quant_fp8's, so the fused kernel
    equals the two-launch form bit for bit when both take the same split-K plan (same fp32 summation order)."""
    from conftest import mk_fp8_weight

    g = torch.Generator().manual_seed(M + K)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16().to(DEV)
    Wq, ws = mk_fp8_weight(N, K)
    Wd, wsd = Wq.to(DEV), ws.to(DEV).view(-1).contiguous()
    out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    with capi.fused_decode():
        capi.check(capi.lib.b2_linear_w8a8_fp8(x.data_ptr(), capi.DTYPE_BF16, K, Wd.data_ptr(), K, wsd.data_ptr(), None,
                                               0, out.data_ptr(), N, M, N, K, None, 0, _stream()))
    cfg = capi.last_gemm_config()
    assert cfg.endswith("_fusedq")
    q, s = b2.sglang_per_token_quant_fp8(x)
    two = b2.fp8_scaled_mm(q, Wd.t(), s, wsd.view(-1, 1), torch.bfloat16, None)
    if capi.last_gemm_config() + "_fusedq" == cfg:
        assert torch.equal(bf16_bits(out.cpu()), bf16_bits(two.cpu()))
    q_ref, s_ref = restate.per_token_quant_fp8(x.cpu())
    ref = restate.fp8_scaled_mm(q_ref, Wq.view(torch.uint8), s_ref, ws, None)
    o = out.float().cpu()
    assert torch.all((o - ref).abs() <= 1.01 * _ulp_bf16(ref) + 1e-6 * ref.abs().max())


# ====================================================== host-buffer pipeline ====

@pytest.mark.parametrize("M,chunk", [(1, 64), (130, 64), (1000, 256), (2048, 1024)])
def test_host_buffer_pipeline_matches_device_path_bitwise(M, chunk):
    """HostW8A8Int8Linear (the e2e entry bench.py times): host activations in, host outputs out, chunked through
    three streams with double-buffered staging -- ragged last chunk, single-row batch, two layers in flight at once
    (submit / wait) -- equals the device-resident op bit for bit."""
    layers = []
    for i, (K, N) in enumerate(((1024, 1536), (2048, 512))):
        x, W, Bs, bias = mk_int8_problem(M, N, K, seed=M + i)
        hl = b2.HostW8A8Int8Linear(W, Bs, bias, device=DEV, chunk_rows=chunk)
        out_host = torch.empty(M, N, dtype=torch.bfloat16).pin_memory()
        layers.append((hl, x.pin_memory(), out_host, W, Bs, bias))
    for _ in range(2):  # second round reuses the staging buffers
        for hl, xh, oh, *_ in layers:
            hl.submit(xh, oh)
        for hl, *_ in layers:
            hl.wait()
    for hl, xh, oh, W, Bs, bias in layers:
        want = b2.int8_scaled_mm_with_quant(xh.to(DEV), W.to(DEV), Bs.to(DEV), bias.to(DEV), torch.bfloat16, False)
        assert torch.equal(oh.view(torch.int16), want.cpu().view(torch.int16))
        assert hl(xh, oh) is oh  # the blocking form
        assert torch.equal(oh.view(torch.int16), want.cpu().view(torch.int16))
