"""Host time per EAGER C-ABI call (no CUDA graph): b2_gemm_s8 at a decode shape, launched back to back through ctypes.
(profiles/r02q_host_overhead.log: ~6 us of host time per call, of which the three cuTensorMapEncodeTiled calls are
below the noise -- a cache of encoded maps was tried and removed.)"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
from sglang_diffusion_b200 import capi
dev = torch.device("cuda:0")
for M, N, K in ((16, 6144, 4096), (8192, 6144, 4096)):
    A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev)
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
    As, Bs = torch.rand(M, device=dev), torch.rand(N, device=dev)
    D = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    args = (A.data_ptr(), K, W.data_ptr(), K, As.data_ptr(), Bs.data_ptr(), None, 0, D.data_ptr(), N, capi.DTYPE_BF16, M, N, K, st)
    for _ in range(20): capi.lib.b2_gemm_s8(*args)
    torch.cuda.synchronize()
    n = 2000 if M == 16 else 200
    t0 = time.perf_counter()
    for _ in range(n): capi.lib.b2_gemm_s8(*args)
    t_issue = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    t_total = (time.perf_counter() - t0) / n
    print(f"b2_gemm_s8 M={M} N={N} K={K} [{capi.last_gemm_config()}]: "
          f"host issue {t_issue*1e6:.2f} us/call, wall {t_total*1e6:.2f} us/call", flush=True)
