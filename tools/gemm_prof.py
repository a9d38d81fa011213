"""Where does the GEMM wait?  Runs one shape per tile config with the in-kernel wait-cycle counters on
(b2_set_profile_buffer) and prints per-role waiting as a fraction of the CTA lifetime.
Usage: python tools/gemm_prof.py [kind] [M N K] [cfgs...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2

kind = sys.argv[1] if len(sys.argv) > 1 else "s8"
M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (8192, 28672, 4096)
cfgs = [int(v) for v in sys.argv[5:]] or [4, 5]
dev = "cuda:0"
zeros = os.environ.get("B2_PROF_DATA") == "zeros"  # zero operands draw little power: structure without the power cap
if kind == "s8":
    A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev)
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
    if zeros:
        A.zero_(); W.zero_()
    op = b2.int8_scaled_mm
else:
    A = torch.randn(M, K, device=dev).to(torch.float8_e4m3fn)
    W = torch.randn(N, K, device=dev).to(torch.float8_e4m3fn)
    if zeros:
        A = torch.zeros(M, K, device=dev).to(torch.float8_e4m3fn); W = torch.zeros(N, K, device=dev).to(torch.float8_e4m3fn)
    op = b2.fp8_scaled_mm
As = torch.rand(M, 1, device=dev)
Bs = torch.rand(N, 1, device=dev)
if kind == "bw":  # block-wise: SFA [M, K/128] M-major, SFB [K/128, N/128] K-major; cfg 1 = 1-CTA, 2 = 2-CTA
    As = torch.rand(K // 128, M, device=dev).t()
    Bs = torch.rand(N // 128, K // 128, device=dev).t()
    op = lambda a, w, sa, sb, dt, bias: b2.fp8_blockwise_scaled_mm(a, w, sa, sb, dt)
prof = torch.zeros(148 * 8 + 64 * 8, dtype=torch.int64, device=dev)
for cfg in cfgs:
    b2.capi.set_gemm_config(cfg)
    for _ in range(3):
        op(A, W.t(), As, Bs, torch.bfloat16, None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        op(A, W.t(), As, Bs, torch.bfloat16, None)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 10 * 1e-3
    b2.capi.lib.b2_set_profile_buffer(prof.data_ptr())
    prof.zero_()
    op(A, W.t(), As, Bs, torch.bfloat16, None)
    torch.cuda.synchronize()
    b2.capi.lib.b2_set_profile_buffer(None)
    tl = prof[148 * 8:].view(64, 8).cpu()
    p = prof[:148 * 8].view(148, 8).double().cpu()
    cfgname = b2.capi.last_gemm_config()
    tm = 256 if cfgname.startswith('2cta') else 128
    tn = int(cfgname.split('x')[1].split('_')[0])
    tiles_total = (-(-M // tm)) * (-(-N // tn)) * (2 if tm == 256 else 1)
    life = p[:, 4].clamp_min(1)
    lead = p[p[:, 0] > 0] if (p[:, 0] > 0).any() else p
    print(f"{kind} M={M} N={N} K={K} cfg={cfg}:{b2.capi.last_gemm_config()}  {t*1e6:8.1f} us {2.0*M*N*K/t/1e12:7.1f} TOPS | "
          f"CTA life {life.mean():.0f} clk (max {life.max():.0f}) | MMA waits operands {100*(lead[:,0]/lead[:,4]).mean():.1f}% "
          f"accum {100*(lead[:,1]/lead[:,4]).mean():.1f}% | producer waits stage {100*(p[:,2]/life).mean():.1f}% | "
          f"epilogue waits accum {100*(p[:,3]/life).mean():.1f}% body {100*(p[:,7]/life).mean():.1f}% "
          f"(store-buffer wait {100*(p[:,5]/life).mean():.1f}%, tmem-ld wait {100*(p[:,6]/life).mean():.1f}%) "
          f"body clk/tile {(p[:,7].sum()/ (tiles_total)):.0f}", flush=True)
    if kind == "bw":
        nkb = K // 128
        kbs = tiles_total / 148 * nkb  # K blocks per CTA
        print(f"   block-wise, clk per K block: loop {life.mean()/kbs:.0f} | promotion warp waits: partial {p[:,3].mean()/kbs:.0f} "
              f"scales+barrier {p[:,5].mean()/kbs:.0f} | MMA warp waits (stage + slot) {lead[:,0].mean()/kbs:.0f} | "
              f"producer waits {p[:,2].mean()/kbs:.0f} | MMA issue block {lead[:,1].mean()/kbs:.0f} | promotion: first 2 LDTM+wait "
              f"{p[:,6].mean()/kbs:.0f} body after wait {p[:,7].mean()/kbs:.0f}", flush=True)
b2.capi.set_gemm_config(0)
