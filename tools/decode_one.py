"""ncu target: one decode linear (quantize + cluster split-K GEMM) per BASELINE C2 layer at M = 16, plus the
quantizers at M = 8192.  python tools/decode_one.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2

dev = torch.device("cuda:0")
M = 16
with b2.capi.static_weights():
    for K, N in ((4096, 6144), (4096, 28672), (14336, 4096)):
        x = (torch.randn(M, K, device=dev) / 10).bfloat16()
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
        Bs = torch.rand(N, 1, device=dev) * 1e-2
        for _ in range(3):
            q, s = b2.per_token_quant_int8(x)
            o = b2.int8_scaled_mm(q, W.t(), s, Bs, torch.bfloat16, None)
        print(K, N, b2.capi.last_gemm_config())
xb = (torch.randn(8192, 4096, device=dev) / 10).bfloat16()
for _ in range(2):
    b2.per_token_quant_int8(xb)
    b2.sglang_per_token_quant_fp8(xb)
    b2.per_token_group_quant_fp8(xb, 128, column_major_scales=True)
torch.cuda.synchronize()
print("ok")
