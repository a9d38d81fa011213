"""One launch sequence of the 1x128 group quantizer (ncu target): python tools/group_quant_one.py [M K]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2

M, K = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8192, 4096)
x = (torch.randn(M, K, device="cuda") / 10).bfloat16()
for _ in range(3):
    q, s = b2.per_token_group_quant_fp8(x, 128, column_major_scales=True)
torch.cuda.synchronize()
print("ok", q.shape, s.shape)
