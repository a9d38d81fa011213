"""Exhaustive (tile width, split, ring depth) sweep of the cluster split-K decode kernel for given shapes, graph
replayed with operands rotating through > L2 bytes.   python tools/splitk_plan_sweep.py int8|fp8 M N K [M N K ...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2

dev = torch.device("cuda:0")
kind = sys.argv[1]
shapes = [tuple(int(v) for v in sys.argv[i:i + 3]) for i in range(2, len(sys.argv), 3)]
b2.capi.set_weight_prefetch(True)


def timed(fns, min_ms=20.0):
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s_ = torch.cuda.Stream()
    s_.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s_):
        with torch.cuda.graph(g, stream=s_):
            for f in fns:
                f()
    torch.cuda.current_stream().wait_stream(s_)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 4
    while True:
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if ms >= min_ms or reps >= 2048:
            return ms * 1e-3 / (reps * len(fns))
        reps *= 2


for M, N, K in shapes:
    nset = max(2, min(48, int(160e6 // (N * K)) + 1))
    if kind == "int8":
        sets = [(torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev),
                 torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)) for _ in range(nset)]
        op = b2.int8_scaled_mm
    else:
        sets = [(torch.randn(M, K, device=dev).to(torch.float8_e4m3fn),
                 torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)) for _ in range(nset)]
        op = b2.fp8_scaled_mm
    As, Bs = torch.rand(M, 1, device=dev), torch.rand(N, 1, device=dev)
    byts = M * K + N * K + 2 * M * N
    res = []
    cands = [0, 2, 3, 4] + [0x80000 | (bn // 16 - 1) | ((S - 1) << 4) | (cap << 8) | (amc << 13)
                   for bn in (32, 64, 96, 128, 192, 256) for S in (1, 2, 4, 8) for cap in (4, 8, 24)
                   for amc in ((0, 1, 2) if M >= 32 else (0,)) if S << amc <= 8]
    for cfg in cands:
        b2.capi.set_gemm_config(cfg)
        try:
            t = timed([(lambda A=A, W=W: op(A, W.t(), As, Bs, torch.bfloat16, None)) for A, W in sets])
            res.append((t, b2.capi.last_gemm_config(), cfg))
        except Exception as e:  # noqa: BLE001
            pass
        finally:
            b2.capi.set_gemm_config(0)
    auto = res[0]
    res.sort()
    print(f"# {kind} M={M} N={N} K={K}: planner {auto[1]} {auto[0]*1e6:.1f} us = {byts/auto[0]/1e9:.0f} GB/s", flush=True)
    for t, name, cfg in res[:10]:
        print(f"   {name:28s} {t*1e6:7.1f} us  {byts/t/1e9:7.0f} GB/s", flush=True)
