"""Forced tile-config sweep of the persistent GEMM on given shapes (graph-replayed, rotating operands).
python tools/cfg_sweep.py int8|fp8 M N K [M N K ...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
dev = torch.device("cuda:0")
kind = sys.argv[1]
shapes = [tuple(int(v) for v in sys.argv[i:i + 3]) for i in range(2, len(sys.argv), 3)]
for M, N, K in shapes:
    nset = max(2, min(16, int(200e6 // (N * K + M * K + 2 * M * N)) + 1))
    if kind == "int8":
        sets = [(torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev), torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)) for _ in range(nset)]
        op = b2.int8_scaled_mm
    else:
        sets = [(torch.randn(M, K, device=dev).to(torch.float8_e4m3fn), torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)) for _ in range(nset)]
        op = b2.fp8_scaled_mm
    As, Bs = torch.rand(M, 1, device=dev), torch.rand(N, 1, device=dev)
    for cfg in ((0, 1, 2, 3, 4) if M <= 128 else (0, 3, 4, 5, 6)):
        b2.capi.set_gemm_config(cfg)
        try:
            fns = [(lambda A=A, W=W: op(A, W.t(), As, Bs, torch.bfloat16, None)) for A, W in sets]
            for f in fns: f()
            name = b2.capi.last_gemm_config()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream(); st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                with torch.cuda.graph(g, stream=st):
                    for f in fns: f()
            torch.cuda.current_stream().wait_stream(st)
            g.replay(); torch.cuda.synchronize()
            best = 1e9
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10): g.replay()
                e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / 10 / nset * 1e3)
        finally:
            b2.capi.set_gemm_config(0)
        print(f"{kind} M={M:<6d} N={N:<6d} K={K:<6d} cfg={cfg} {name:16s} {best:8.1f} us {2.0*M*N*K/best/1e6:8.1f} TOPS", flush=True)
