"""Where does the power go?  Sustained throughput of the int8 GEMM (bench data) with parts of the kernel switched
off through the developer flags (upper bits of b2_set_gemm_config): 4 = no operand loads (the MMAs keep chewing the
real tiles the previous normal launch left in shared memory), 2 = no epilogue body.  Under the 1000 W cap the
throughput gained by removing a part measures the energy that part costs."""
import os, sys, threading, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
import pynvml
M, N, K = 8192, 28672, 4096
dev = "cuda:0"
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
A = (torch.randn(M, K, device=dev) * 32).round().clamp(-127, 127).to(torch.int8)
W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
As = torch.rand(M, 1, device=dev); Bs = torch.rand(N, 1, device=dev)
def sample(stop, out):
    while not stop.is_set():
        out.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)); time.sleep(0.01)
for cfg in (5, 4):
    for dbg, name in ((0, "full kernel"), (2, "no epilogue"), (4, "no loads"), (6, "no loads, no epilogue")):
        b2.capi.set_gemm_config(cfg)
        for _ in range(3):
            b2.int8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, None)   # real tiles end up resident in smem
        b2.capi.set_gemm_config(cfg | (dbg << 8))
        torch.cuda.synchronize()
        stop, samples = threading.Event(), []
        th = threading.Thread(target=sample, args=(stop, samples)); th.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 1500
        e0.record()
        for _ in range(n):
            b2.int8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, None)
        e1.record(); torch.cuda.synchronize(); stop.set(); th.join()
        tail = samples[len(samples) // 2:]
        clk = sorted(s[0] for s in tail)[len(tail) // 2]; pw = max(s[1] for s in tail)
        print(f"cfg {b2.capi.last_gemm_config():14s} {name:24s} sustained {2.0*M*N*K*n/(e0.elapsed_time(e1)*1e-3)/1e12:7.1f} TOPS | SM clock {clk} MHz, power max {pw:.0f} W", flush=True)
b2.capi.set_gemm_config(0)
