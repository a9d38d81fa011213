"""Library yardstick for ncu: torch._int_mm (cuBLASLt s8 x s8 -> s32) on the bench shape and data."""
import torch
M, N, K = 8192, 28672, 4096
A = (torch.randn(M, K, device="cuda") * 32).round().clamp(-127, 127).to(torch.int8)
W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device="cuda")
for _ in range(4):
    torch._int_mm(A, W.t())
torch.cuda.synchronize()
