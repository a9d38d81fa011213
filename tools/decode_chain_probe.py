"""Decode step (3 Llama-3-8B layers at M = 16, quantizer + GEMM each), graph-replayed: plain launches vs the
static-weight chain.  B2_QUANT_PDL=0 keeps the quantizer's launch plain (A/B of the quantizer chaining)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2

dev = torch.device("cuda:0")
M = 16
layers = [(4096, 6144), (4096, 28672), (14336, 4096)]
probs = []
for K, N in layers:
    probs.append(((torch.randn(M, K, device=dev) / 10).bfloat16(),
                  torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev), torch.rand(N, 1, device=dev) * 1e-2))


def step():
    for x, W, Bs in probs:
        q, s = b2.per_token_quant_int8(x)
        b2.int8_scaled_mm(q, W.t(), s, Bs, torch.bfloat16, None)


def timed(static):
    b2.capi.set_weight_prefetch(static)
    step()
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            step()
    torch.cuda.current_stream().wait_stream(st)
    b2.capi.set_weight_prefetch(False)
    for _ in range(10):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record()
        for _ in range(200):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / 200)
    return best


byts = sum(2 * M * K + N * K + 2 * M * N + 4 * N for K, N in layers)
for name, static in (("plain launches", False), ("static-weight chain", True)):
    us = timed(static)
    print(f"{name:22s} B2_QUANT_PDL={os.environ.get('B2_QUANT_PDL', '1')}  {us:7.2f} us/step  {byts / us / 1e3:7.1f} GB/s", flush=True)
