"""Per-shape throughput of every kernel family on the BASELINE.json configs (developer tool; feeds profiles/).

    python tools/perf_sweep.py [int8] [fp8] [block] [quant] [--quick]

Method: CUDA events around back-to-back launches for >= 40 ms after warm-up; operands rotate through
enough distinct buffers that the bytes touched between two uses of the same buffer exceed the 126 MB L2, so the
memory-bound shapes (M <= 64, quantizers) see HBM and not a warm cache.  Ops = 2*M*N*K; bytes = the ALGORITHMIC
bytes of SURVEY 8(d) (every operand once, output once).  Peaks: tensor = 2 x the measured cuBLAS bf16 rate of
MEASURED_PEAKS.json (burst, these are short runs), HBM = its measured copy bandwidth.
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sglang_diffusion_b200 as b2  # noqa: E402

dev = torch.device("cuda:0")
QUICK = "--quick" in sys.argv
want = [a for a in sys.argv[1:] if not a.startswith("--")] or ["int8", "fp8", "block", "quant"]
try:
    PK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    PK = {}
HBM = PK.get("hbm_gbs", 6650.0)
TENSOR = 2.0 * PK.get("bf16_tflops", 1590.0)
L2_BYTES = 160e6


PDL = "--pdl" in sys.argv  # declare the weights static: decode GEMMs launch programmatically (b2_set_weight_prefetch)
b2.capi.set_weight_prefetch(PDL)
GRAPH = "--graph" in sys.argv  # replay the launches from a CUDA graph: kernel time without the Python/dispatcher cost


def timed(fns, min_ms=40.0):
    """fns: list of closures (one per rotating buffer set).  Returns seconds per call."""
    for f in fns[: min(len(fns), 3)]:
        f()
    torch.cuda.synchronize()
    if GRAPH:
        g = torch.cuda.CUDAGraph()
        s_ = torch.cuda.Stream()
        s_.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s_):
            for f in fns:
                f()
            with torch.cuda.graph(g, stream=s_):
                for f in fns:
                    f()
        torch.cuda.current_stream().wait_stream(s_)
        torch.cuda.synchronize()
        inner = fns
        fns = [g.replay]
        n_inner = len(inner)
    else:
        n_inner = 1
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = len(fns)
    reps = 1
    while True:
        e0.record()
        for _ in range(reps):
            for f in fns:
                f()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if ms >= min_ms or reps >= 4096:
            return ms * 1e-3 / (reps * n * n_inner)
        reps = max(reps * 2, int(reps * min_ms / max(ms, 1e-3)) + 1)


def copies_for(bytes_per_set):
    return max(1, min(64, int(L2_BYTES // max(bytes_per_set, 1)) + 1))


def line(name, M, N, K, t, ops, byts, cfg=""):
    tops = ops / t / 1e12
    gbs = byts / t / 1e9
    print(f"{name:22s} M={M:<6d} N={N:<6d} K={K:<6d} {t*1e6:9.1f} us  {tops:8.1f} TOPS ({100*tops/TENSOR:5.1f}% of {TENSOR:.0f})  "
          f"{gbs:8.1f} GB/s ({100*gbs/HBM:5.1f}% of {HBM:.0f})  {cfg}", flush=True)


def gemm_int8(M, N, K, bias=False, cfg=0):
    nset = copies_for(N * K + M * K + 2 * M * N)
    sets = []
    for _ in range(nset):
        A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev)
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
        sets.append((A, W))
    As = torch.rand(M, 1, device=dev)
    Bs = torch.rand(N, 1, device=dev) * 1e-2
    b = torch.randn(N, device=dev).bfloat16() if bias else None
    fns = [(lambda A=A, W=W: b2.int8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, b)) for A, W in sets]
    b2.capi.set_gemm_config(cfg)
    try:
        t = timed(fns)
    finally:
        b2.capi.set_gemm_config(0)
    line("int8_scaled_mm", M, N, K, t, 2.0 * M * N * K, M * K + N * K + 2 * M * N + 4 * M + 4 * N, b2.capi.last_gemm_config())


def gemm_fp8(M, N, K):
    nset = copies_for(N * K + M * K + 2 * M * N)
    sets = []
    for _ in range(nset):
        A = (torch.randn(M, K, device=dev)).to(torch.float8_e4m3fn)
        W = (torch.randn(N, K, device=dev)).to(torch.float8_e4m3fn)
        sets.append((A, W))
    As = torch.rand(M, 1, device=dev)
    Bs = torch.rand(N, 1, device=dev) * 1e-2
    fns = [(lambda A=A, W=W: b2.fp8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, None)) for A, W in sets]
    t = timed(fns)
    line("fp8_scaled_mm", M, N, K, t, 2.0 * M * N * K, M * K + N * K + 2 * M * N + 4 * M + 4 * N, b2.capi.last_gemm_config())


def gemm_block(M, N, K):
    nset = copies_for(N * K + M * K + 2 * M * N)
    sets = []
    for _ in range(nset):
        A = (torch.randn(M, K, device=dev)).to(torch.float8_e4m3fn)
        W = (torch.randn(N, K, device=dev)).to(torch.float8_e4m3fn)
        sets.append((A, W))
    sfa = torch.rand(K // 128, M, device=dev).t()          # [M, K/128], stride(0) == 1 (M-major)
    sfb = torch.rand(N // 128, K // 128, device=dev).t()   # [K/128, N/128], K-major
    fns = [(lambda A=A, W=W: b2.fp8_blockwise_scaled_mm(A, W.t(), sfa, sfb, torch.bfloat16)) for A, W in sets]
    t = timed(fns)
    line("fp8_blockwise_mm", M, N, K, t, 2.0 * M * N * K,
         M * K + N * K + 2 * M * N + 4 * M * K // 128 + 4 * (N // 128) * (K // 128), b2.capi.last_gemm_config())


def quant(kind, M, K):
    nset = copies_for(3 * M * K)
    xs = [(torch.randn(M, K, device=dev) / 10).bfloat16() for _ in range(nset)]
    if kind == "int8":
        fns = [(lambda x=x: b2.per_token_quant_int8(x)) for x in xs]
        byts = 3 * M * K + 4 * M
    elif kind == "fp8":
        fns = [(lambda x=x: b2.sglang_per_token_quant_fp8(x)) for x in xs]
        byts = 3 * M * K + 4 * M
    else:
        fns = [(lambda x=x: b2.sglang_per_token_group_quant_fp8(x, 128, column_major_scales=True)) for x in xs]
        byts = 3 * M * K + 4 * M * K // 128
    t = timed(fns)
    line(f"quant_{kind}", M, 0, K, t, 0.0, byts)


def linear_int8(M, N, K):
    # quantizer + GEMM per call, as W8A8Int8LinearMethod.apply issues them
    nset = copies_for(N * K + 2 * M * K + 2 * M * N)
    sets = []
    for _ in range(nset):
        x = (torch.randn(M, K, device=dev) / 10).bfloat16()
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
        sets.append((x, W))
    Bs = torch.rand(N, device=dev) * 1e-2
    fns = [(lambda x=x, W=W: b2.int8_scaled_mm_with_quant(x, W, Bs, None, torch.bfloat16, False)) for x, W in sets]
    t = timed(fns)
    line("int8 linear (q+mm)", M, N, K, t, 2.0 * M * N * K, 2 * M * K + N * K + 2 * M * N + 4 * N)


def fused(kind, mode, M, K):
    """Producer-fused quantizers against the unfused pair (torch eager producer + separate quantizer).  Bytes =
    algorithmic bytes of the FUSED op: silu 2*2K in + K (+scales) out per row; rmsnorm with residual 2K + 2K in,
    2K (residual) + K out."""
    if kind == "silu":
        nset = copies_for(5 * M * K)
        xs = [torch.randn(M, 2 * K, device=dev).bfloat16() for _ in range(nset)]
        fns = [(lambda x=x: b2.silu_and_mul_quant(x, mode)) for x in xs]
        byts = 4 * M * K + M * K + (4 * M if "group" not in mode else 4 * M * K // 128)
    else:
        nset = copies_for(7 * M * K)
        xs = [(torch.randn(M, K, device=dev).bfloat16(), torch.randn(M, K, device=dev).bfloat16()) for _ in range(nset)]
        w = torch.ones(K, device=dev).bfloat16()
        fns = [(lambda x=x, r=r: b2.rmsnorm_quant(x, w, 1e-6, residual=r, mode=mode)) for x, r in xs]
        byts = 6 * M * K + M * K + (4 * M if "group" not in mode else 4 * M * K // 128)
    t = timed(fns)
    line(f"{kind}->{mode}", M, 0, K, t, 0.0, byts)


if "fused" in want:
    print("# producer-fused quantizers (bytes = algorithmic bytes of the fused op)")
    for M, K in ((8192, 14336), (8192, 4096), (32768, 5120), (4096, 2048), (16, 14336)):
        for mode in ("int8", "fp8", "fp8_group128_v2"):
            fused("silu", mode, M, K)
        for mode in ("int8", "fp8", "fp8_group128"):
            fused("rms", mode, M, K)
    for M, K in ((8192, 4096), (8192, 14336)):
        x = (torch.randn(M, K, device=dev) / 10).bfloat16()
        nset = copies_for(3 * M * K)
        xs = [(torch.randn(M, K, device=dev) / 10).bfloat16() for _ in range(nset)]
        st = torch.tensor([0.01], device=dev)
        t = timed([(lambda x=x: b2.scaled_fp8_quant(x, st)) for x in xs])
        line("per_tensor static", M, 0, K, t, 0.0, 3 * M * K)
        t = timed([(lambda x=x: b2.scaled_fp8_quant(x)) for x in xs])
        line("per_tensor dynamic", M, 0, K, t, 0.0, 5 * M * K)
        t = timed([(lambda x=x: b2.sglang_per_token_group_quant_fp8(x, 128, column_major_scales=True, enable_v2=True)) for x in xs])
        line("group128 v2", M, 0, K, t, 0.0, 3 * M * K + 4 * M * K // 128)
def moe(T, E, topk, H, I):
    """fused_experts (W8A8-int8 MoE): six launches; TOPS over the routed MACs 2 * T * topk * 3 * I * H."""
    g = torch.Generator(device=dev).manual_seed(T + E)
    x = (torch.randn(T, H, generator=g, device=dev) / 10).bfloat16()
    w13 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8, generator=g, device=dev)
    w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8, generator=g, device=dev)
    s13 = torch.rand(E, 2 * I, 1, generator=g, device=dev) * 2e-3
    s2 = torch.rand(E, H, 1, generator=g, device=dev) * 2e-3
    tw, ids = torch.topk(torch.softmax(torch.randn(T, E, generator=g, device=dev), -1), topk, dim=-1)
    ids = ids.to(torch.int32)
    t = timed([lambda: b2.fused_experts(x, w13, w2, tw, ids, s13, s2)])
    ops = 2.0 * T * topk * 3 * I * H
    byts = E * 3 * I * H + 2 * T * H * 2  # every expert weight once + activations in / out
    tops = ops / t / 1e12
    print(f"fused_experts int8     T={T:<6d} E={E:<4d} topk={topk} H={H:<5d} I={I:<5d} {t*1e6:9.1f} us  {tops:8.1f} TOPS "
          f"({100*tops/TENSOR:5.1f}% of {TENSOR:.0f})  {byts/t/1e9:8.1f} GB/s of weights+io", flush=True)


if "bmm" in want:
    print("# bmm_fp8 (MLA absorbed BMMs, per-tensor scales): one batched launch; cuBLAS bf16 bmm of the dequantized operands beside it")
    for b, m, k, n in ((128, 16, 128, 512), (128, 16, 512, 128), (16, 4096, 128, 512), (16, 4096, 512, 128), (128, 256, 128, 512)):
        A = torch.randn(b, m, k, device=dev).to(torch.float8_e4m3fn)
        W = torch.randn(b, n, k, device=dev).to(torch.float8_e4m3fn)
        sa, sb = torch.tensor([0.01], device=dev), torch.tensor([0.02], device=dev)
        t = timed([lambda: b2.bmm_fp8(A, W.transpose(-2, -1), sa, sb, torch.bfloat16)])
        Ab, Wb = A.bfloat16(), W.bfloat16().transpose(-2, -1)
        t2 = timed([lambda: torch.bmm(Ab, Wb)])
        print(f"bmm_fp8 b={b:<4d} m={m:<5d} k={k:<4d} n={n:<4d} {t*1e6:8.1f} us  {2.0*b*m*n*k/t/1e12:7.1f} TFLOPS   "
              f"[{b2.capi.last_gemm_config()}]   torch.bmm bf16 {t2*1e6:8.1f} us", flush=True)
    sys.exit(0)
if "moe" in want:
    print("# W8A8-int8 MoE (fused_experts): Qwen1.5-MoE-A2.7B-like (E=60->64, topk 4, H 2048, I 1408), DeepSeek-V2-Lite-like (E=64, topk 6)")
    for T in (16, 256, 4096, 16384):
        moe(T, 64, 4, 2048, 1408)
        moe(T, 64, 6, 2048, 1408)
    moe(4096, 8, 2, 4096, 14336)  # Mixtral-8x7B
    sys.exit(0)
if "splitk" in want:
    # decode kernel (gemm_splitk_sm100.cu): the planner's pick, then every forced (BN, S), per C2 decode shape
    for K, N in ((4096, 6144), (4096, 28672), (14336, 4096), (8192, 1280), (1024, 8192)):
        for M in ((16, 128) if not QUICK else (16,)):
            print(f"# M={M} N={N} K={K}")
            gemm_int8(M, N, K)
            gemm_int8(M, N, K, cfg=1 if N <= 6144 else 4)
            if "--all" in sys.argv:
                for bn in (32, 64, 96, 112, 128, 176, 208, 256):
                    for S in (1, 2, 4, 8):
                        for cap in (4, 24):
                            gemm_int8(M, N, K, cfg=0x80000 | (bn // 16 - 1) | ((S - 1) << 4) | (cap << 8))
    sys.exit(0)
if "int8" in want:
    print("# C1 + C2: Llama-3-8B int8 shapes"); gemm_int8(128, 4096, 4096, bias=True)
    Ms = [1, 16, 64, 256, 1024, 4096, 8192] if not QUICK else [16, 8192]
    if "--small" in sys.argv:
        Ms = [1, 16, 64, 128]
    for K, N in ((4096, 6144), (4096, 28672), (14336, 4096)):
        for M in Ms:
            gemm_int8(M, N, K)
    if "--small" in sys.argv:
        sys.exit(0)
    print("# C4: Wan2.1-14B / FLUX.1 DiT shapes, M = 32768 and the 8-way shard 4096")
    for K, N in ((5120, 5120), (5120, 13824), (13824, 5120), (3072, 9216), (3072, 12288), (12288, 3072)):
        for M in ([32768, 4096] if not QUICK else [4096]):
            gemm_int8(M, N, K)
    print("# fused linear"); linear_int8(8192, 28672, 4096); linear_int8(16, 28672, 4096)
if "fp8" in want:
    print("# C5: Llama-3-70B TP=8 fp8 shards")
    for K, N in ((8192, 1280), (8192, 7168), (1024, 8192), (3584, 8192)):
        for M in (64, 8192):
            gemm_fp8(M, N, K)
    gemm_fp8(8192, 28672, 4096)
if "block" in want:
    print("# C3: DeepSeek-V3 block-wise fp8, M = 4096")
    for N, K in ((24576, 1536), (32768, 512), (7168, 16384), (36864, 7168), (7168, 18432), (4096, 7168), (7168, 2048)):
        gemm_block(4096, N, K)
if "quant" in want:
    print("# quantizers")
    for M, K in ((8192, 4096), (8192, 14336), (32768, 5120), (64, 8192), (16, 4096)):
        for kind in ("int8", "fp8", "group"):
            quant(kind, M, K)
