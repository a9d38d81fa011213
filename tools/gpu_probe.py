"""First-contact GPU diagnostics (developer tool, not a test): runs every kernel family on a few
shapes, prints mismatch statistics instead of asserting, and times the GEMM tile configs.
Usage (on a B200): python tools/gpu_probe.py [--quick]
"""
from __future__ import annotations

import os
import sys
import time
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import sglang_diffusion_b200 as b2  # noqa: E402
from oracle import restate  # noqa: E402  (checker only)

dev = torch.device("cuda:0")
QUICK = "--quick" in sys.argv


def section(name):
    print(f"\n==== {name} ====", flush=True)


def guarded(fn):
    def run(*a, **k):
        try:
            return fn(*a, **k)
        except Exception:
            traceback.print_exc()
            print("!! FAILED", fn.__name__, a, flush=True)
            try:
                torch.cuda.synchronize()
            except Exception as e:  # sticky error: nothing else will work
                print("!! CUDA context is dead:", e, flush=True)
                sys.exit(3)
    return run


def mk_int8(M, N, K, seed=1234):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    W = ((torch.rand(N, K, generator=g) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
    Bs = torch.rand(N, generator=g) * 1e-2
    bias = torch.randn(N, generator=g).bfloat16()
    return x, W, Bs, bias


@guarded
def probe_quant_int8(M, K):
    x, _, _, _ = mk_int8(M, 16, K)
    q_ref, s_ref = restate.per_token_quant_int8_cpu(x)
    q, s = b2.per_token_quant_int8_cpu(x.to(dev))
    torch.cuda.synchronize()
    q = q.cpu()
    s = s.cpu()
    nq = (q.to(torch.int16) + 128 != q_ref.to(torch.int16)).sum().item()
    ns = (s.view(torch.int32) != s_ref.view(torch.int32)).sum().item()
    print(f"quant_int8 M={M} K={K}: code mismatches {nq}/{q.numel()} scale mismatches {ns}/{M}", flush=True)


@guarded
def probe_gemm_s8_acc(M, N, K, cfg):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g)
    W = torch.randint(-128, 128, (N, K), dtype=torch.int8, generator=g)
    ref = (A.double() @ W.double().t()).to(torch.int32)  # exact (|acc| < 2^53)
    Ad, Wd = A.to(dev), W.to(dev)
    D = torch.full((M, N), -7, dtype=torch.int32, device=dev)
    b2.capi.set_gemm_config(cfg)
    st = b2.capi.lib.b2_gemm_s8(Ad.data_ptr(), K, Wd.data_ptr(), K, None, None, None, 0, D.data_ptr(), N,
                                b2.capi.DTYPE_S32, M, N, K, torch.cuda.current_stream().cuda_stream)
    b2.capi.set_gemm_config(0)
    if st != 0:
        print(f"gemm_s8 acc M={M} N={N} K={K} cfg={cfg}: status {st} {b2.capi.last_error()}")
        return
    torch.cuda.synchronize()
    got = D.cpu()
    bad = (got != ref).sum().item()
    print(f"gemm_s8 acc M={M} N={N} K={K} cfg={cfg} ({b2.capi.last_gemm_config()}): mismatches {bad}/{ref.numel()}",
          flush=True)
    if bad:
        idx = (got != ref).nonzero()
        rows = idx[:, 0].unique()
        cols = idx[:, 1].unique()
        print(f"   bad rows: {rows[:16].tolist()}{'...' if len(rows) > 16 else ''} ({len(rows)})  "
              f"bad cols: {cols[:16].tolist()}{'...' if len(cols) > 16 else ''} ({len(cols)})")
        i, j = idx[0].tolist()
        print(f"   first: [{i},{j}] got {got[i, j].item()} want {ref[i, j].item()}")


@guarded
def probe_linear_int8(M, N, K, has_bias, fused):
    x, W, Bs, bias = mk_int8(M, N, K)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, bias.float() if has_bias else None)
    xd, Wd, Bsd, bd = x.to(dev), W.to(dev), Bs.to(dev), bias.to(dev)
    if fused:
        out = b2.int8_scaled_mm_with_quant(xd, Wd, Bsd, bd if has_bias else None, torch.bfloat16, False)
    else:
        q, s = b2.per_token_quant_int8(xd)
        out = b2.int8_scaled_mm(q, Wd.t(), s, Bsd.view(-1, 1), torch.bfloat16, bd if has_bias else None)
    torch.cuda.synchronize()
    bad = (out.cpu().view(torch.int16) != ref.view(torch.int16)).sum().item()
    print(f"int8 linear M={M} N={N} K={K} bias={has_bias} fused={fused}: bf16 mismatches {bad}/{ref.numel()}",
          flush=True)


def mk_fp8(M, N, K, seed=7):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    w = torch.randn(N, K, generator=g) * 0.05
    ws = w.abs().amax(dim=1, keepdim=True) / 448.0
    Wq = (w / ws).clamp(-448, 448).to(torch.float8_e4m3fn)
    bias = torch.randn(N, generator=g).bfloat16()
    return x, Wq, ws.float(), bias


@guarded
def probe_quant_fp8(M, K):
    x, _, _, _ = mk_fp8(M, 16, K)
    q_ref, s_ref = restate.per_token_quant_fp8(x)
    q, s = b2.sglang_per_token_quant_fp8(x.to(dev))
    torch.cuda.synchronize()
    nq = (q.cpu().view(torch.uint8) != q_ref).sum().item()
    ns = (s.cpu().view(-1).view(torch.int32) != s_ref.view(torch.int32)).sum().item()
    print(f"quant_fp8 M={M} K={K}: code mismatches {nq}/{q.numel()} scale mismatches {ns}/{M}", flush=True)


@guarded
def probe_quant_group(M, K):
    x, _, _, _ = mk_fp8(M, 16, K)
    q_ref, s_ref = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    q, s = b2.per_token_group_quant_fp8(x.to(dev), 128, column_major_scales=True)
    torch.cuda.synchronize()
    nq = (q.cpu().view(torch.uint8) != q_ref).sum().item()
    ns = (s.cpu().contiguous().view(torch.int32) != s_ref.contiguous().view(torch.int32)).sum().item()
    print(f"quant_group128 M={M} K={K}: code mismatches {nq}/{q.numel()} scale mismatches {ns}/{s.numel()} "
          f"scale strides {tuple(s.stride())}", flush=True)


@guarded
def probe_gemm_fp8(M, N, K, has_bias, cfg=0):
    x, Wq, ws, bias = mk_fp8(M, N, K)
    q_ref, s_ref = restate.per_token_quant_fp8(x)
    ref = restate.fp8_scaled_mm(q_ref, Wq.view(torch.uint8), s_ref, ws, bias.float() if has_bias else None)
    qd = q_ref.to(dev).view(torch.float8_e4m3fn)
    b2.capi.set_gemm_config(cfg)
    out = b2.fp8_scaled_mm(qd, Wq.to(dev).t(), s_ref.to(dev).view(-1, 1), ws.to(dev), torch.bfloat16,
                           bias.to(dev) if has_bias else None)
    b2.capi.set_gemm_config(0)
    torch.cuda.synchronize()
    o = out.float().cpu()
    err = (o - ref).abs()
    print(f"fp8 gemm M={M} N={N} K={K} bias={has_bias} cfg={b2.capi.last_gemm_config()}: max abs err {err.max():.4g} "
          f"max|ref| {ref.abs().max():.4g} normalised {err.max() / ref.abs().max():.3g} "
          f"bf16-exact {(o == ref.bfloat16().float()).float().mean():.4f}", flush=True)


@guarded
def probe_gemm_block(M, N, K):
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    w = torch.randn(N, K, generator=g) * 0.05
    wb = w.view(N // 128, 128, K // 128, 128)
    sfb = wb.abs().amax(dim=(1, 3)) / 448.0
    Wq = (wb / sfb[:, None, :, None]).clamp(-448, 448).reshape(N, K).to(torch.float8_e4m3fn)
    q_ref, sfa = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
    ref = restate.fp8_blockwise_scaled_mm(q_ref, Wq.view(torch.uint8), sfa, sfb)
    qd = q_ref.to(dev).view(torch.float8_e4m3fn)
    sfa_d = sfa.t().contiguous().to(dev).t()  # M-major
    sfb_d = sfb.to(dev)  # [N/128, K/128] row-major; op wants [K/128, N/128] with stride(0)==1 -> .t()
    out = b2.fp8_blockwise_scaled_mm(qd, Wq.to(dev).t(), sfa_d, sfb_d.t(), torch.bfloat16)
    torch.cuda.synchronize()
    o = out.float().cpu()
    err = (o - ref).abs()
    print(f"fp8 block gemm M={M} N={N} K={K}: max abs err {err.max():.4g} max|ref| {ref.abs().max():.4g} "
          f"normalised {err.max() / ref.abs().max():.3g}", flush=True)


def time_cuda(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


@guarded
def bench_gemm(kind, M, N, K, cfg):
    if kind == "s8":
        A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev)
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
        op = b2.int8_scaled_mm
    else:
        A = torch.randn(M, K, device=dev).to(torch.float8_e4m3fn)
        W = torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)
        op = b2.fp8_scaled_mm
    As = torch.rand(M, 1, device=dev)
    Bs = torch.rand(N, 1, device=dev)
    b2.capi.set_gemm_config(cfg)
    t = time_cuda(lambda: op(A, W.t(), As, Bs, torch.bfloat16, None))
    name = b2.capi.last_gemm_config()
    b2.capi.set_gemm_config(0)
    tops = 2.0 * M * N * K / t / 1e12
    gbs = (M * K + N * K + 2 * M * N) / t / 1e9
    print(f"bench {kind} M={M} N={N} K={K} cfg={cfg}:{name}  {t * 1e6:9.1f} us  {tops:8.1f} TOPS  {gbs:8.1f} GB/s",
          flush=True)


@guarded
def bench_quant(M, K):
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    t = time_cuda(lambda: b2.per_token_quant_int8_cpu(x))
    print(f"bench quant_int8 M={M} K={K}: {t * 1e6:.1f} us {(3.0 * M * K + 4 * M) / t / 1e9:.0f} GB/s", flush=True)
    xq = torch.empty(M, K, device=dev, dtype=torch.float8_e4m3fn)
    xs = torch.empty(M, 1, device=dev, dtype=torch.float32)
    t = time_cuda(lambda: b2.sgl_per_token_quant_fp8(x, xq, xs))
    print(f"bench quant_fp8  M={M} K={K}: {t * 1e6:.1f} us {(3.0 * M * K + 4 * M) / t / 1e9:.0f} GB/s", flush=True)


def main():
    print(torch.cuda.get_device_name(0), b2.capi.device_info(), flush=True)
    t0 = time.time()
    section("int8 quantizer parity")
    for M, K in [(1, 512), (3, 544), (128, 4096), (257, 14336), (7, 16384)]:
        probe_quant_int8(M, K)
    section("int8 GEMM raw accumulators (exact)")
    for cfg in (3, 4, 2, 1, 5, 6):
        probe_gemm_s8_acc(128, 256, 256, cfg)
    for cfg in (1, 2, 3, 4, 5, 6):
        probe_gemm_s8_acc(300, 1000, 544, cfg)
    probe_gemm_s8_acc(2048, 2048, 1024, 5)
    probe_gemm_s8_acc(2048, 2048, 1024, 6)
    probe_gemm_s8_acc(1, 4096, 4096, 0)
    probe_gemm_s8_acc(5, 384, 544, 0)
    probe_gemm_s8_acc(1000, 6144, 4096, 0)
    section("int8 linear bf16 (bit-exact vs oracle)")
    for M, N, K in [(2, 384, 544), (128, 384, 544), (128, 4096, 4096)]:
        for hb in (False, True):
            for fused in (False, True):
                probe_linear_int8(M, N, K, hb, fused)
    section("fp8 quantizers")
    for M, K in [(1, 512), (128, 1076 // 4 * 4), (128, 4096), (300, 8192)]:
        probe_quant_fp8(M, K)
    for M, K in [(1, 512), (127, 7168), (512, 2048)]:
        probe_quant_group(M, K)
    section("fp8 GEMMs")
    for M, N, K in [(64, 1280, 8192), (300, 7168, 1024), (5, 128, 512)]:
        probe_gemm_fp8(M, N, K, True)
        probe_gemm_fp8(M, N, K, False)
    for M, N, K in [(1, 512, 512), (127, 2048, 7168), (512, 4096, 1536)]:
        probe_gemm_block(M, N, K)
    print(f"\nparity probes took {time.time() - t0:.1f}s", flush=True)
    if QUICK:
        return
    section("timing")
    for M, K in [(128, 4096), (8192, 4096), (32768, 5120), (16, 14336)]:
        bench_quant(M, K)
    for kind in ("s8", "f8"):
        for cfg in (3, 4, 5, 6):
            bench_gemm(kind, 8192, 6144, 4096, cfg)
        for cfg in (4, 5):
            bench_gemm(kind, 8192, 28672, 4096, cfg)
            bench_gemm(kind, 8192, 4096, 14336, cfg)
        bench_gemm(kind, 128, 4096, 4096, 0)
        for cfg in (1, 2, 3):
            bench_gemm(kind, 16, 28672, 4096, cfg)
        bench_gemm(kind, 1, 4096, 14336, 0)


if __name__ == "__main__":
    main()
