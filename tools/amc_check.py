import sys, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
dev = "cuda:0"
torch.manual_seed(0)
bad = 0
for kind in ("int8", "fp8"):
    for M, N, K in ((64, 7168, 8192), (64, 1024, 1024), (128, 4096, 4096), (32, 2048, 2048), (40, 1536, 640), (64, 8192, 3584), (96, 512, 4096)):
        if kind == "int8":
            A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev); W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
            op = b2.int8_scaled_mm
        else:
            A = torch.randn(M, K, device=dev).to(torch.float8_e4m3fn); W = torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)
            op = b2.fp8_scaled_mm
        As, Bs = torch.rand(M, 1, device=dev), torch.rand(N, 1, device=dev)
        bias = torch.randn(N, device=dev).bfloat16()
        for bn in (32, 64, 128):
            for S in (1, 2, 4):
                base = 0x80000 | (bn // 16 - 1) | ((S - 1) << 4)
                b2.capi.set_gemm_config(base)
                ref = op(A, W.t(), As, Bs, torch.bfloat16, bias); n0 = b2.capi.last_gemm_config()
                for amc in (1, 2):
                    if S << amc > 8: continue
                    b2.capi.set_gemm_config(base | (amc << 13))
                    out = op(A, W.t(), As, Bs, torch.bfloat16, bias); n1 = b2.capi.last_gemm_config()
                    torch.cuda.synchronize()
                    if not torch.equal(out.view(torch.int16), ref.view(torch.int16)):
                        bad += 1
                        print("MISMATCH", kind, M, N, K, n0, n1, (out.view(torch.int16) != ref.view(torch.int16)).sum().item(), flush=True)
                b2.capi.set_gemm_config(0)
print("AMC_CHECK", "PASS" if bad == 0 else f"FAIL {bad}", flush=True)
