"""One launch sequence of a producer-fused quantizer (ncu target): python tools/fused_quant_one.py silu|rms mode M K"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
kind, mode, M, K = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
if kind == "silu":
    x = torch.randn(M, 2 * K, device="cuda").bfloat16()
    for _ in range(3):
        b2.silu_and_mul_quant(x, mode)
else:
    x = torch.randn(M, K, device="cuda").bfloat16(); r = torch.randn(M, K, device="cuda").bfloat16()
    w = torch.ones(K, device="cuda").bfloat16()
    for _ in range(3):
        b2.rmsnorm_quant(x, w, 1e-6, residual=r, mode=mode)
torch.cuda.synchronize()
print("ok")
