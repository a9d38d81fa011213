"""Is the int8 GEMM limited by its structure or by the board's power cap?  Runs the same shape with operand
data of different toggle activity (zeros, +-1, small gaussian, full-range random) for ~1.5 s each and prints the
throughput together with the SM clock / power NVML reports while it runs.
Usage: python tools/power_probe.py [kind] [M N K] [cfg]"""
import os, sys, threading, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2
import pynvml

kind = sys.argv[1] if len(sys.argv) > 1 else "s8"
M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (8192, 28672, 4096)
cfg = int(sys.argv[5]) if len(sys.argv) > 5 else 0
dev = "cuda:0"
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)


def sample(stop, out):
    while not stop.is_set():
        out.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
        time.sleep(0.01)


def make(mode):
    if mode == "zeros":
        A = torch.zeros(M, K, dtype=torch.int8, device=dev); W = torch.zeros(N, K, dtype=torch.int8, device=dev)
    elif mode == "pm1":
        A = (torch.randint(0, 2, (M, K), device=dev) * 2 - 1).to(torch.int8)
        W = (torch.randint(0, 2, (N, K), device=dev) * 2 - 1).to(torch.int8)
    elif mode == "gauss32":
        A = (torch.randn(M, K, device=dev) * 32).round().clamp(-127, 127).to(torch.int8)
        W = (torch.randn(N, K, device=dev) * 32).round().clamp(-127, 127).to(torch.int8)
    elif mode == "bench":  # bench.py's distribution: activations ~N(0, 32), weights uniform
        A = (torch.randn(M, K, device=dev) * 32).round().clamp(-127, 127).to(torch.int8)
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
    else:
        A = torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev)
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)
    if kind == "f8":
        A = A.view(torch.uint8).bitwise_and(0x77).view(torch.float8_e4m3fn)
        W = W.view(torch.uint8).bitwise_and(0x77).view(torch.float8_e4m3fn)
    return A, W


op = b2.int8_scaled_mm if kind == "s8" else b2.fp8_scaled_mm
As = torch.rand(M, 1, device=dev)
Bs = torch.rand(N, 1, device=dev)
b2.capi.set_gemm_config(cfg)
for mode in os.environ.get("B2_PROBE_MODES", "zeros,pm1,gauss32,bench,uniform").split(","):
    A, W = make(mode)
    for _ in range(3):
        op(A, W.t(), As, Bs, torch.bfloat16, None)
    torch.cuda.synchronize()
    # burst: 5 launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    time.sleep(0.5)
    e0.record()
    for _ in range(5):
        op(A, W.t(), As, Bs, torch.bfloat16, None)
    e1.record(); torch.cuda.synchronize()
    burst = 2.0 * M * N * K * 5 / (e0.elapsed_time(e1) * 1e-3) / 1e12
    stop, samples = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, samples)); th.start()
    n = 1500
    e0.record()
    for _ in range(n):
        op(A, W.t(), As, Bs, torch.bfloat16, None)
    e1.record(); torch.cuda.synchronize()
    stop.set(); th.join()
    sus = 2.0 * M * N * K * n / (e0.elapsed_time(e1) * 1e-3) / 1e12
    tail = samples[len(samples) // 2:]
    clk = sorted(s[0] for s in tail)[len(tail) // 2] if tail else 0
    pw = max(s[1] for s in tail) if tail else 0
    print(f"{kind} {mode:8s} cfg={b2.capi.last_gemm_config()} burst {burst:7.1f} TOPS | sustained {sus:7.1f} TOPS over "
          f"{e0.elapsed_time(e1):.0f} ms | SM clock median {clk} MHz, power max {pw:.0f} W", flush=True)
b2.capi.set_gemm_config(0)

# library yardsticks on the same shape / data (not the reference; cuBLASLt through torch)
if os.environ.get("B2_PROBE_MODES"):
    sys.exit(0)
if kind == "s8":
    A, W = make("bench")
    Wt = W.t()
    for name, fn in (("torch._int_mm (s8->s32, no epilogue)", lambda: torch._int_mm(A, Wt)),):
        try:
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(300):
                fn()
            e1.record(); torch.cuda.synchronize()
            print(f"{name}: {2.0*M*N*K*300/(e0.elapsed_time(e1)*1e-3)/1e12:.1f} TOPS sustained", flush=True)
        except Exception as ex:  # noqa: BLE001
            print(f"{name}: unavailable ({type(ex).__name__}: {ex})")
else:
    A, W = make("bench")
    sa = torch.rand(M, 1, device=dev); sb = torch.rand(1, N, device=dev)
    try:
        fn = lambda: torch._scaled_mm(A, W.t(), scale_a=sa, scale_b=sb, out_dtype=torch.bfloat16)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(300):
            fn()
        e1.record(); torch.cuda.synchronize()
        print(f"torch._scaled_mm rowwise fp8: {2.0*M*N*K*300/(e0.elapsed_time(e1)*1e-3)/1e12:.1f} TFLOPS sustained", flush=True)
    except Exception as ex:  # noqa: BLE001
        print(f"torch._scaled_mm: unavailable ({type(ex).__name__}: {ex})")
