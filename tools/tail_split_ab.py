"""A/B of the persistent GEMM's tail split (half-width work items for a ragged last wave): time + bit-exactness.
python tools/tail_split_ab.py int8|fp8 M N K [M N K ...]      (debug flag 8 in the forced-config word turns the split off)"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
dev = torch.device("cuda:0")
kind = sys.argv[1]
shapes = [tuple(int(v) for v in sys.argv[i:i + 3]) for i in range(2, len(sys.argv), 3)]
torch.manual_seed(0)
for M, N, K in shapes:
    nset = max(2, min(16, int(200e6 // (N * K + M * K + 2 * M * N)) + 1))
    if kind == "int8":
        sets = [(torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev), torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)) for _ in range(nset)]
        op = b2.int8_scaled_mm
    else:
        sets = [(torch.randn(M, K, device=dev).to(torch.float8_e4m3fn), torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)) for _ in range(nset)]
        op = b2.fp8_scaled_mm
    As, Bs = torch.rand(M, 1, device=dev), torch.rand(N, 1, device=dev)
    bias = torch.randn(N, device=dev, dtype=torch.bfloat16)
    res, outs, graphs = {}, {}, {}
    for name, word in (("split", 0), ("nosplit", 8 << 8)):
        b2.capi.set_gemm_config(word)
        try:
            fns = [(lambda A=A, W=W: op(A, W.t(), As, Bs, torch.bfloat16, bias)) for A, W in sets]
            outs[name] = fns[0]().clone()
            cfg = b2.capi.last_gemm_config()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream(); st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                with torch.cuda.graph(g, stream=st):
                    for f in fns: f()
            torch.cuda.current_stream().wait_stream(st)
            g.replay(); torch.cuda.synchronize()
            graphs[name] = g
            res[name] = 1e9
        finally:
            b2.capi.set_gemm_config(0)
    for _ in range(6):  # interleaved, so clock drift under the power cap hits both arms alike
        for name in ("nosplit", "split"):
            g = graphs[name]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): g.replay()
            e1.record(); torch.cuda.synchronize()
            res[name] = min(res[name], e0.elapsed_time(e1) / 10 / nset * 1e3)
    same = torch.equal(outs["split"], outs["nosplit"])
    print(f"{kind} M={M} N={N} K={K} cfg={cfg}: split {res['split']:.1f} us ({2*M*N*K/res['split']/1e6:.0f} T)  "
          f"nosplit {res['nosplit']:.1f} us ({2*M*N*K/res['nosplit']/1e6:.0f} T)  ratio {res['nosplit']/res['split']:.3f}  bit-equal {same}", flush=True)
