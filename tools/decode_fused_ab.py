"""Per-layer A/B of the one-kernel decode linear (FUSEQ) against quantizer + GEMM, graph replayed with rotating weights.
Run twice: B2_FUSED_DECODE=1 (default) and B2_FUSED_DECODE=0."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
dev = torch.device("cuda:0")
b2.capi.set_weight_prefetch(True)
for M in (1, 16, 32):
    for K, N in ((4096, 6144), (4096, 28672), (14336, 4096), (8192, 1280), (1024, 8192)):
        nset = max(2, int(160e6 // (N * K)) + 1)
        Ws = [torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev) for _ in range(nset)]
        x = (torch.randn(M, K, device=dev) / 10).bfloat16()
        Bs = torch.rand(N, device=dev) * 1e-2
        fns = [(lambda W=W: b2.int8_scaled_mm_with_quant(x, W, Bs, None, torch.bfloat16, False)) for W in Ws]
        for f in fns: f()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream(); st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            with torch.cuda.graph(g, stream=st):
                for f in fns: f()
        torch.cuda.current_stream().wait_stream(st)
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): g.replay()
        e1.record(); torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / 20 / nset * 1e3
        print(f"M={M:<3d} K={K:<6d} N={N:<6d} {t:7.2f} us  {(N*K+2*M*K+2*M*N)/t/1e3:7.0f} GB/s  {b2.capi.last_gemm_config()}", flush=True)
