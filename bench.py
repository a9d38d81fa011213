"""bench.py -- the W8A8 quantized-linear hot path on B200, one JSON line per run.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A "step" = one pass of the hot path (dynamic per-token int8 quantization -> tcgen05 int8 GEMM -> scale/bias
epilogue -> bf16) over one batch of M synthetic tokens through every linear of the workload.  Default workload
(BASELINE.json configs[1]): the Llama-3-8B W8A8-int8 linears QKV 4096->6144, gate_up 4096->28672,
down 14336->4096 at M = 8192 tokens per GPU.  Weights are random-init int8 with per-channel fp32 scales and stay
resident in HBM; activations are synthetic bf16 (reference test distribution, test/srt/cpu/test_gemm.py:133).

  value     TOPS (2*M*N*K summed over the layers / device time), inputs already in HBM, CUDA events, max over ranks
  e2e       same metric through the host-buffer entry (HostW8A8Int8Linear): pinned host activations -> device,
            quantize + GEMM, bf16 outputs -> pinned host, every step, copies inside the timed region
  roofline  the int8 GEMM kernel alone (events around its launches): achieved TOPS vs 2 x the measured cuBLAS bf16
            peak of MEASURED_PEAKS.json (int8 tensor rate = 2 x bf16 on sm_100a); datasheet fraction alongside
  cpu_baseline  the reference's own CPU path (oracle/_ref, compiled from /root/reference) -- or the oracle port
            where that cannot run -- on the host cores, on a bounded row sample of the same workload
  --impl reference   times only that CPU path (rank 0), same metric/config keys

N > 1: one process per GPU, every rank runs the full workload on its own M tokens (rows are independent:
M-sharded data parallel, weights replicated, NO collective on the data path) -> "scaling": "weak".  The same line
then also carries (the rows the DP headline cannot show):
  strong            BASELINE configs[3]: the Wan2.1-14B DiT linears on M = 32768 / N rows per GPU (strong scaling)
  tp                BASELINE configs[4]: Llama-3-70B W8A8-fp8 decoder-layer linears at TP = N (qkv, o_proj, gate_up,
                    down; column / row shards), decode M = 64 and prefill M = 8192, row-parallel sums once through
                    NCCL and once through b2_allreduce_symm (NVSwitch peer memory)
  multi_gpu_parity  ordered-sum bit-equality of the peer-memory all-reduce, TP linears against fp64 shard products,
                    M-sharded rows against the unsharded result; any failure makes the run exit non-zero
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (description, M per GPU, [(layer, K, N)], has_bias)
    "llama3-8b-w8a8-int8": ("Llama-3-8B W8A8-int8 linears (qkv 4096x6144, gate_up 4096x28672, down 14336x4096), "
                            "M=8192 tokens/GPU", 8192,
                            [("qkv_proj", 4096, 6144), ("gate_up_proj", 4096, 28672), ("down_proj", 14336, 4096)], False),
    "c1-int8-128x4096x4096": ("int8 W8A8 linear M=128 K=4096 N=4096 bf16 out, bias", 128, [("linear", 4096, 4096)], True),
    "wan2.1-14b-dit-int8": ("Wan2.1-14B DiT W8A8-int8 linears (5120x5120, 5120x13824, 13824x5120), M=32768 tokens "
                            "sharded over the GPUs", 32768,
                            [("attn_proj", 5120, 5120), ("ffn_up", 5120, 13824), ("ffn_down", 13824, 5120)], True),
}
DEFAULT_WORKLOAD = "llama3-8b-w8a8-int8"
DATASHEET_INT8_TOPS = 4500.0
CPU_SAMPLE_ROWS = 1024


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per GEMM launch (average over the workload's layers) from the
    committed `ncu --set full` capture, or None when no capture of this workload is on file."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return d["avg_bytes_per_launch"] if d.get("workload") == workload else None
    except (OSError, ValueError, KeyError):
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_burst": d.get("bf16_tflops", 1590.0),
                "bf16_sustained": d.get("bf16_tflops_sustained", 1400.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


def mma_only_peak():
    """tcgen05.mma issue-only peak of this pool's B200s (tools/mma_peak.cu: TMEM accumulators, operands resident in
    shared memory, no loads, no epilogue), recorded in profiles/mma_only_peak.json from profiles/r01a_mma_only_peak.log."""
    try:
        with open(os.path.join(ROOT, "profiles", "mma_only_peak.json")) as f:
            d = json.load(f)
        return {k: d[k] for k in ("int8_burst", "int8_sustained", "fp8_burst", "fp8_sustained", "source")}
    except (OSError, ValueError, KeyError):
        return {"int8_burst": 4559.0, "int8_sustained": 4059.0, "fp8_burst": 4360.0, "fp8_sustained": 3758.0,
                "source": "fallback constants (profiles/mma_only_peak.json missing)"}


def total_ops(M, layers):
    return sum(2.0 * M * N * K for _, K, N in layers)


# ------------------------------------------------------------------ clocks ----

class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, device_index: int, period_s: float = 0.002):
        super().__init__(daemon=True)
        self.period = period_s
        self.samples, self.reason_bits, self.power = [], 0, []
        self.sm_max = None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        while not self._stop_evt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.reason_bits |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(self.period)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        s = sorted(self.samples)
        reasons = [n for b, n in self.REASONS.items() if self.reason_bits & b and n != "gpu_idle"]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.sm_max, "reasons": reasons, "samples": len(s),
                "power_w_max": round(max(self.power), 1) if self.power else None}


# ------------------------------------------------------------ CPU reference ----

def cpu_reference_run(layers, has_bias, sample_rows, warmup, steps, min_seconds=0.0):
    """Times the reference CPU path on the host cores: int8_scaled_mm_with_quant on `sample_rows` rows of
    every layer of the workload (weights prepacked once, untimed, as the reference does at load).
    Returns (TOPS, seconds per step, kind, cores, steps actually timed)."""
    from oracle import build_ref, restate

    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    use_ref = build_ref.load()
    kind = "reference" if use_ref else "port"
    rows = sample_rows if use_ref else min(sample_rows, 64)  # the scalar port is ~100x slower
    g = torch.Generator().manual_seed(1234)
    probs = []
    for _, K, N in layers:
        x = (torch.randn(rows, K, generator=g) / 10).bfloat16()
        W = ((torch.rand(N, K, generator=g) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
        Bs = torch.rand(N, generator=g) * 1e-2
        b = torch.randn(N, generator=g).bfloat16().float() if has_bias else None
        if use_ref:
            W = torch.ops.sgl_kernel_ref.convert_weight_packed(W)
        probs.append((x, W, Bs, b))

    def step():
        for x, W, Bs, b in probs:
            if use_ref:
                torch.ops.sgl_kernel_ref.int8_scaled_mm_with_quant(x, W, Bs, b, torch.bfloat16, True)
            else:
                restate.int8_scaled_mm_with_quant(x, W, Bs, b)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    n = 0
    while n < steps or (time.perf_counter() - t0) < min_seconds:
        step()
        n += 1
        if n >= steps and min_seconds == 0.0:
            break
        if n >= 100000:
            break
    dt = (time.perf_counter() - t0) / n
    tops = total_ops(rows, layers) / dt / 1e12
    sample = (f"{rows} token rows of each of the {len(layers)} layers per step, {n} steps, weights prepacked once "
              f"(untimed); {'oracle/_ref (reference sources compiled in place, AMX/AVX512)' if use_ref else 'oracle/restate.c scalar port (host lacks AMX/AVX512-BF16 or oracle/_ref missing)'}")
    return tops, dt, kind, cores, n, sample, rows


# ------------------------------------------------------------------- main ----

def decode_companion(b2, dev, layers, probs, peaks):
    """north_star: ">= 80 % of HBM bandwidth for M <= 16".  The workload's layers at M = 16 through the linear method's
    call sequence (quantizer -> cluster split-K GEMM, weights declared static), replayed from a CUDA graph.  HBM-bound
    on the weight bytes; the layers' weights (> L2 in total) rotate, so every replay streams them from HBM.
    Informational: not part of `value`; measured BEFORE the sustained large-M loop (a power-capped board right after
    it runs its fabric slower: 58 us instead of 49 us per step)."""
    try:
        Md = 16
        xs = [(torch.randn(Md, K, generator=torch.Generator(device=dev).manual_seed(7), device=dev) / 10).bfloat16()
              for _, K, _ in layers]

        def decode_step():
            # the linear method's call at decode M: ONE op per layer (row quantizer + cluster split-K GEMM, chained
            # programmatically; the one-kernel form b2_set_fused_decode is measured slower and stays off)
            with b2.capi.static_weights():
                for xd, (x, W, Bs, b) in zip(xs, probs):
                    b2.int8_scaled_mm_with_quant(xd, W, Bs, b, torch.bfloat16, False)

        decode_step()
        torch.cuda.synchronize(dev)
        gst = torch.cuda.Stream(dev)
        gst.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(gst):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=gst):
                decode_step()
        torch.cuda.current_stream(dev).wait_stream(gst)
        for _ in range(5):
            gr.replay()
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nrep = 100
        d0.record()
        for _ in range(nrep):
            gr.replay()
        d1.record()
        torch.cuda.synchronize(dev)
        dec_us = d0.elapsed_time(d1) * 1e3 / nrep
        dec_bytes = sum(2 * Md * K + N * K + 2 * Md * N + 4 * N for _, K, N in layers)
        dec_gbs = dec_bytes / (dec_us * 1e-6) / 1e9
        decode = {"workload": f"same layers at M = {Md}: int8_scaled_mm_with_quant per layer (row quantizer + cluster "
                              "split-K GEMM, programmatically chained), graph replay",
                  "bound": "hbm", "us_per_step": round(dec_us, 2), "achieved": round(dec_gbs, 1),
                  "peak": round(peaks["hbm_gbs"], 1), "unit": "GB/s", "frac": round(dec_gbs / peaks["hbm_gbs"], 4),
                  "algorithmic_bytes": dec_bytes, "gemm_config_last_layer": b2.capi.last_gemm_config()}
    except Exception as e:  # noqa: BLE001  (informational leg: never take the headline line down with it)
        decode = {"error": f"{type(e).__name__}: {e}"}
    return decode



# ------------------------------------------------------------ multi-GPU legs ----

WAN_LAYERS = WORKLOADS["wan2.1-14b-dit-int8"][2]
TP_LAYERS = [  # Llama-3-70B, full sizes: name, parallel, K, N  (bench_fp8_gemm.py:47-52)
    ("qkv_proj", "column", 8192, 10240), ("o_proj", "row", 8192, 8192),
    ("gate_up_proj", "column", 8192, 57344), ("down_proj", "row", 28672, 8192)]


def _max_over_ranks(ms, dev, dist, world):
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _time_device(fn, dev, dist, world, iters, warm=3, graph_inner=0):
    """Device time per call of fn (CUDA events on the current stream, barrier + synchronize on both sides, MAX over
    ranks).  graph_inner > 0: `graph_inner` calls are captured into one CUDA graph and the replays are timed."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    run, per = fn, 1
    if graph_inner:
        st = torch.cuda.Stream(dev)
        st.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(st):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for _ in range(graph_inner):
                    fn()
        torch.cuda.current_stream(dev).wait_stream(st)
        g.replay()
        torch.cuda.synchronize(dev)
        run, per = g.replay, graph_inner
    reps = max(1, iters // per)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize(dev)
    return _max_over_ranks(e0.elapsed_time(e1) / (reps * per), dev, dist, world)


def strong_scaling_leg(b2, dev, dist, world, rank, steps):
    """BASELINE configs[3]: Wan2.1-14B DiT W8A8-int8 linears, 32768 diffusion tokens split over the GPUs."""
    M = 32768 // world
    g = torch.Generator(device=dev).manual_seed(99 + rank)
    probs = []
    for _, K, N in WAN_LAYERS:
        x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
        Bs = torch.rand(N, generator=g, device=dev) * 1e-2
        b = torch.randn(N, generator=g, device=dev).bfloat16()
        probs.append((x, W, Bs, b))

    def step():
        for x, W, Bs, b in probs:
            b2.int8_scaled_mm_with_quant(x, W, Bs, b, torch.bfloat16, False)

    ms = _time_device(step, dev, dist, world, max(5, min(steps, 20)))
    ops_total = total_ops(32768, WAN_LAYERS)
    return {"workload": "Wan2.1-14B DiT W8A8-int8 linears (5120x5120, 5120x13824, 13824x5120), M = 32768 total",
            "scaling": "strong", "rows_per_gpu": M, "ms_per_step": round(ms, 4),
            "value": round(ops_total / (ms * 1e-3) / 1e12, 1), "unit": "TOPS",
            "tokens_per_s": round(32768 / (ms * 1e-3), 1)}


def tp_leg(b2, dev, dist, world, rank, comm):
    """BASELINE configs[4]: the four linears of a Llama-3-70B decoder layer, W8A8-fp8, TP = world."""
    from sglang_diffusion_b200.linear import QuantLinear, W8A8Fp8LinearMethod

    method = W8A8Fp8LinearMethod()
    g = torch.Generator(device=dev).manual_seed(4321 + rank)
    layers = []
    for name, par, K, N in TP_LAYERS:
        layer = QuantLinear(K, N, method, bias=False, parallel=par, tp_rank=rank, tp_size=world, device=dev)
        n_loc, k_loc = (N // world, K) if par == "column" else (N, K // world)
        w = torch.randn(n_loc, k_loc, generator=g, device=dev) * 0.05
        ws = w.abs().amax(dim=1, keepdim=True) / 448.0
        layer.weight.data.copy_((w / ws).to(torch.float8_e4m3fn))
        layer.weight_scale.data.copy_(ws)
        method.process_weights_after_loading(layer)
        del w
        layers.append((name, par, K, N, k_loc, layer))
    rows, check = [], {"max_rel_err": 0.0, "ok": True}
    for M in (16, 64, 8192):
        xs = [(torch.randn(M, k_loc, generator=g, device=dev) / 10).bfloat16() for *_, k_loc, _ in layers]

        def step():
            for x, (*_, layer) in zip(xs, layers):
                layer(x)

        res = {"M": M, "tp": world}
        for mode in ("nccl", "b2_allreduce_symm", "b2_fused_push", "b2_fused_scatter_allgather"):
            for _, par, _, _, _, layer in layers:
                layer.symm_comm = comm if (mode != "nccl" and par == "row") else None
                layer.symm_fused_push = mode == "b2_fused_push"
                layer.symm_fused_scatter = mode == "b2_fused_scatter_allgather"
                layer.symm_force = mode != "nccl"  # measure the named path even where the layer's own policy prefers NCCL
            if mode == "b2_fused_scatter_allgather" and (comm is None or not comm.fits_scatter_all_reduce(M, 8192, torch.bfloat16)):
                res[mode] = None  # scatter epilogue + gathered sum + all-gather: M % (32 * world) == 0
                continue
            if mode != "nccl" and (comm is None or not comm.fits(M, 8192, torch.bfloat16)):
                res[mode] = None
                continue
            if mode == "b2_fused_push" and not comm.fits_push(M, 8192, torch.bfloat16):
                res[mode] = None  # GEMM-epilogue push is built for decode shapes (M <= 128) with a multicast mapping
                continue
            ms = _time_device(step, dev, dist, world, 200 if M <= 64 else 10, graph_inner=10 if M <= 64 else 0)
            per_layer = {}
            for x, (name, par, K, N, _, layer) in zip(xs, layers):
                per_layer[name] = round(_time_device(lambda: layer(x), dev, dist, world, 100 if M <= 64 else 5,
                                                     graph_inner=10 if M <= 64 else 0) * 1e3, 1)
            ops = sum(2.0 * M * N * K for _, _, K, N, _, _ in layers)
            res[mode] = {"us_per_decoder_layer_linears": round(ms * 1e3, 1), "tops_aggregate": round(ops / (ms * 1e-3) / 1e12, 1),
                         "us_per_layer": per_layer}
            # parity of the row-parallel layers against the fp64 product of the same quantized shards (M = 64 only:
            # the fp64 reference of the prefill shape would dominate the run)
            if M == 64:
                for x, (name, par, K, N, _, layer) in zip(xs, layers):
                    if par != "row":
                        continue
                    out = layer(x)
                    wdq = layer.weight.t().double() * layer.weight_scale.double().view(-1, 1)
                    part = x.double() @ wdq.t()
                    dist.all_reduce(part)
                    err = ((out.double() - part).abs().max() / part.abs().max()).item()
                    check["max_rel_err"] = max(check["max_rel_err"], err)
                    # dynamic e4m3 activations vs the exact product of the same weights: 3 significand bits per element
                    check["ok"] = check["ok"] and err < 6e-2
        for *_, layer in layers:
            layer.symm_comm = None
            layer.symm_fused_push = None
            layer.symm_fused_scatter = None
            layer.symm_force = False
        if M >= 32 * world and M % (32 * world) == 0:
            # prefill, row-sharded sum (RowParallelLinear(skip_all_reduce) + LayerCommunicator._tp_reduce_scatter of the
            # reference): GEMM + NCCL reduce_scatter_tensor against the GEMM's scatter epilogue + one local sum
            def step_rs():
                for x, (_, par, _, _, _, layer) in zip(xs, layers):
                    layer.forward_reduce_scatter(x) if par == "row" else layer(x)

            for mode in ("nccl_reduce_scatter", "b2_fused_reduce_scatter"):
                fused = mode == "b2_fused_reduce_scatter"
                if fused and (comm is None or not comm.fits_scatter(M, 8192, torch.bfloat16)):
                    res[mode] = None
                    continue
                for _, par, _, _, _, layer in layers:
                    layer.symm_comm = comm if (fused and par == "row") else None
                ms = _time_device(step_rs, dev, dist, world, 10)
                per_layer = {}
                for x, (name, par, K, N, _, layer) in zip(xs, layers):
                    if par == "row":
                        per_layer[name] = round(_time_device(lambda: layer.forward_reduce_scatter(x), dev, dist, world, 5) * 1e3, 1)
                res[mode] = {"us_per_decoder_layer_linears": round(ms * 1e3, 1), "us_per_row_layer": per_layer}
                if fused:
                    # against the NCCL reduce-scatter of the same partials (different summation order: one bf16 ulp)
                    for x, (name, par, K, N, _, layer) in zip(xs, layers):
                        if par != "row":
                            continue
                        got = layer.forward_reduce_scatter(x).float()
                        layer.symm_comm = None
                        want = layer.forward_reduce_scatter(x).float()
                        layer.symm_comm = comm
                        err = ((got - want).abs().max() / want.abs().max()).item()
                        check["fused_reduce_scatter_vs_nccl_max_rel_err"] = max(check.get("fused_reduce_scatter_vs_nccl_max_rel_err", 0.0), round(err, 5))
                        check["ok"] = check["ok"] and err < 1.6e-2
            for *_, layer in layers:
                layer.symm_comm = None
            res["nvlink_bytes_out_per_rank_per_row_layer_reduce_scatter"] = (world - 1) * M * 8192 * 2 // world
        # row-parallel layer -> all-reduce -> fused_add_rmsnorm -> quantizer of the next layer (what a decoder block does
        # behind o_proj / down_proj): NCCL all-reduce + the fused norm/quant kernel against the library's two launches
        # (decode: pushed GEMM + gathered sum/norm/quant; prefill: scatter GEMM + sequence-parallel sum/norm/quant +
        # all-gather of the codes by peer stores)
        try:
            from sglang_diffusion_b200 import ops as _ops
            gn = torch.Generator(device=dev).manual_seed(7)
            wn = (torch.rand(8192, generator=gn, device=dev) + 0.5).bfloat16()
            res_full = torch.randn(M, 8192, generator=gn, device=dev).bfloat16()
            seq_par = M > 128
            can_fuse = comm is not None and M % world == 0 and (
                comm.fits_scatter_rmsnorm_allgather(M, 8192, torch.bfloat16) if seq_par
                else comm.fits_reduce_rmsnorm(M, 8192, torch.bfloat16))
            rows_loc = M // world if M % world == 0 else 0
            res_loc = res_full[rank * rows_loc:(rank + 1) * rows_loc].clone()
            chain = {}
            for mode in ("nccl_allreduce+rmsnorm_quant", "b2_fused_norm_chain"):
                fused = mode == "b2_fused_norm_chain"
                if fused and not can_fuse:
                    chain[mode] = None
                    continue
                per_layer = {}
                for x, (name, par, K, N, _, layer) in zip(xs, layers):
                    if par != "row":
                        continue
                    layer.symm_comm = comm if fused else None
                    if not fused:
                        fn = lambda: _ops.rmsnorm_quant(layer(x), wn, 1e-6, residual=res_full, mode="fp8")  # noqa: E731
                    elif seq_par:
                        fn = lambda: layer.forward_reduce_scatter_rmsnorm_quant(x, wn, 1e-6, residual_local=res_loc, mode="fp8")  # noqa: E731
                    else:
                        fn = lambda: layer.forward_allreduce_rmsnorm_quant(x, wn, 1e-6, residual=res_full, mode="fp8")  # noqa: E731
                    per_layer[name] = round(_time_device(fn, dev, dist, world, 100 if M <= 64 else 5,
                                                         graph_inner=10 if M <= 64 else 0) * 1e3, 1)
                    layer.symm_comm = None
                chain[mode] = {"us_per_row_layer_incl_norm_and_quant": per_layer}
            if can_fuse:  # parity of the fused chain against the NCCL composition (other summation order: a few bf16 ulps)
                for x, (name, par, K, N, _, layer) in zip(xs, layers):
                    if par != "row":
                        continue
                    r1 = res_full.clone()
                    q1, s1 = _ops.rmsnorm_quant(layer(x), wn, 1e-6, residual=r1, mode="fp8")
                    layer.symm_comm = comm
                    if seq_par:
                        q2, s2 = layer.forward_reduce_scatter_rmsnorm_quant(x, wn, 1e-6, residual_local=res_loc.clone(), mode="fp8")
                    else:
                        q2, s2 = layer.forward_allreduce_rmsnorm_quant(x, wn, 1e-6, residual=res_full.clone(), mode="fp8")
                    layer.symm_comm = None
                    d1, d2 = q1.float() * s1, q2.float() * s2
                    err = ((d1 - d2).abs().max() / d1.abs().max()).item()
                    check["fused_norm_chain_vs_nccl_max_rel_err"] = max(check.get("fused_norm_chain_vs_nccl_max_rel_err", 0.0), round(err, 5))
                    check["ok"] = check["ok"] and err < 0.15  # one e4m3 step (2^-3 of the value) where a code sits on a rounding boundary
                    dist.barrier()
            res["norm_chain"] = chain
        except Exception as e:  # noqa: BLE001  (a leg that cannot run must not take the headline down)
            res["norm_chain_error"] = f"{type(e).__name__}: {e}"
            for *_, layer in layers:
                layer.symm_comm = None
        res["allreduce_bytes_per_row_layer"] = M * 8192 * 2
        rows.append(res)
    check["max_rel_err"] = round(check["max_rel_err"], 5)
    del layers
    return rows, check


def symm_allreduce_parity(dev, dist, world, rank, comm):
    """b2_allreduce_symm: algo 1 (ordered peer loads) equals the fp32 rank-order sum rounded once, bit for bit; algo 2
    (NVSwitch multimem) is within 1 ulp of the exact sum; both identical on every rank."""
    out = {"algo1_ordered_sum_bitwise": True, "algo2_within_1ulp": None, "same_on_all_ranks": True, "cases": 0}
    algos = [1] + ([2] if comm.multicast_ptr else [])
    if 2 in algos:
        out["algo2_within_1ulp"] = True
    for dtype in (torch.bfloat16, torch.float16):
        for M, N in ((1, 8), (1, 4096), (64, 8192), (100, 5120), (256, 8192)):
            g = torch.Generator(device=dev).manual_seed(1000 * rank + M + N)
            for algo in algos:
                part = comm.partial(M, N, dtype)
                part.copy_(torch.randn(M, N, generator=g, device=dev).to(dtype))
                parts = [torch.empty_like(part) for _ in range(world)]
                dist.all_gather(parts, part.clone())
                acc = parts[0].float()
                for p in parts[1:]:
                    acc = acc + p.float()
                got = comm.all_reduce(part, algo)
                torch.cuda.synchronize(dev)
                if algo == 1:
                    ok = torch.equal(got.view(torch.int16), acc.to(dtype).view(torch.int16))
                    out["algo1_ordered_sum_bitwise"] = out["algo1_ordered_sum_bitwise"] and ok
                else:
                    ok = bool(((got.float() - acc).abs() <= torch.finfo(dtype).eps * acc.abs().clamp_min(1e-3)).all())
                    out["algo2_within_1ulp"] = out["algo2_within_1ulp"] and ok
                ref0 = got.clone()
                dist.broadcast(ref0, 0)
                out["same_on_all_ranks"] = out["same_on_all_ranks"] and torch.equal(ref0.view(torch.int16), got.view(torch.int16))
                out["cases"] += 1
    flags = torch.tensor([int(out["algo1_ordered_sum_bitwise"]), int(out["algo2_within_1ulp"] is not False),
                          int(out["same_on_all_ranks"])], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    out["algo1_ordered_sum_bitwise"], out["same_on_all_ranks"] = bool(flags[0].item()), bool(flags[2].item())
    if out["algo2_within_1ulp"] is not None:
        out["algo2_within_1ulp"] = bool(flags[1].item())
    out["ok"] = bool(flags.min().item())
    # GEMM-epilogue-fused form: equal, bit for bit, to GEMM + one-shot all-reduce (same rank-ordered fp32 sum)
    if comm.multicast_ptr:
        from sglang_diffusion_b200 import ops

        eq = True
        for M, N, K in ((16, 8192, 1024), (64, 8192, 3584)):
            for it in range(3):
                g = torch.Generator(device=dev).manual_seed(77 * rank + M + it)
                x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
                W = torch.randn(N, K, generator=g, device=dev).to(torch.float8_e4m3fn)
                Bs = torch.rand(N, generator=g, device=dev) * 1e-2
                q, s = ops.sglang_per_token_quant_fp8(x)
                part = ops.scaled_mm_out(q, W.t(), s.view(-1), Bs, comm.partial(M, N, torch.bfloat16), None)
                want = comm.all_reduce(part, 1)
                comm.push_gemm(q, W.t(), s.view(-1), Bs, torch.bfloat16, None)
                got = comm.reduce_gathered(M, N, torch.bfloat16)
                torch.cuda.synchronize(dev)
                eq = eq and torch.equal(got.view(torch.int16), want.view(torch.int16))
        f2 = torch.tensor([int(eq)], device=dev)
        dist.all_reduce(f2, op=dist.ReduceOp.MIN)
        out["fused_push_equals_one_shot_bitwise"] = bool(f2.item())
        out["ok"] = out["ok"] and out["fused_push_equals_one_shot_bitwise"]
    return out


def m_sharded_parity(b2, dev, dist, world, rank):
    """Rows of the M-sharded DP run equal the unsharded result bit for bit (per-token scales: no cross-row term)."""
    M, K, N = 1024, 4096, 6144
    g = torch.Generator(device=dev).manual_seed(2024)  # same problem on every rank
    x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
    Bs = torch.rand(N, generator=g, device=dev) * 1e-2
    r0, r1 = b2.shard_rows(M, rank, world)
    mine = b2.int8_scaled_mm_with_quant(x[r0:r1].contiguous(), W, Bs, None, torch.bfloat16, False)
    full = b2.int8_scaled_mm_with_quant(x, W, Bs, None, torch.bfloat16, False)
    ok = torch.tensor([int(torch.equal(mine.view(torch.int16), full[r0:r1].view(torch.int16)))], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return {"rows_equal_unsharded_bitwise": bool(ok.item()), "ok": bool(ok.item())}


_RESULT_FD = None


def emit(line: dict) -> None:
    """The ONE JSON line of the run, written to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    # Libraries chat on stdout (NCCL prints "NCCL version ..." there at init): keep the original stdout for the
    # result line only and send everything else to stderr, so that stdout carries exactly one JSON line.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-decode", action="store_true", help="skip the M = 16 decode companion (profiling runs)")
    ap.add_argument("--no-multi", action="store_true", help="N > 1: skip the strong / tp / multi_gpu_parity legs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    desc, M_cfg, layers, has_bias = WORKLOADS[args.workload]
    sharded_total = args.workload.startswith("wan")
    M = M_cfg // max(world, 1) if sharded_total else M_cfg
    scaling = "strong" if sharded_total else "weak"
    config = {"workload": desc, "name": args.workload, "tokens_per_gpu": M,
              "layers": [{"name": n, "K": K, "N": N} for n, K, N in layers], "bias": has_bias,
              "parallelism": f"dp{world} (M-sharded rows, weights replicated, no collective)",
              "l2": "weights + activations per step exceed the 126 MB L2 (inputs larger than L2; no flush needed)"
                    if sum(K * N for _, K, N in layers) > 126e6 else "L2 flushed between steps by a 256 MB memset"}

    # ------------------------------------------------------ reference arm ----
    if args.impl == "reference":
        if rank != 0:
            return
        tops, dt, kind, cores, n, sample, rows = cpu_reference_run(layers, has_bias, CPU_SAMPLE_ROWS, args.warmup,
                                                                   args.steps)
        config["tokens_per_gpu"] = rows
        line = {"impl": "reference", "metric": "w8a8_int8_linear_tops", "value": round(tops, 4), "unit": "TOPS",
                "n_gpus": args.gpus, "steps": n, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 4),
                "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int8",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": round(tops, 4), "unit": "TOPS", "cores": cores, "kind": kind, "sample": sample},
                "e2e": {"value": round(tops, 4), "unit": "TOPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return

    # ------------------------------------------------------------ our arm ----
    import torch.distributed as dist

    import sglang_diffusion_b200 as b2  # fails loudly when the CUDA library is missing

    assert torch.cuda.is_available(), "bench.py (impl ours) needs a B200; there is no CPU fallback"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize(dev)

    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    probs = []
    for _, K, N in layers:
        x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
        Bs = torch.rand(N, generator=g, device=dev) * 1e-2
        b = torch.randn(N, generator=g, device=dev).bfloat16() if has_bias else None
        probs.append((x, W, Bs, b))
    flush = None
    if sum(K * N for _, K, N in layers) <= 126e6:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    decode = None if args.no_decode else decode_companion(b2, dev, layers, probs, load_peaks())

    def step():
        outs = []
        for x, W, Bs, b in probs:
            outs.append(b2.int8_scaled_mm_with_quant(x, W, Bs, b, torch.bfloat16, False))
        return outs

    ops_step = total_ops(M, layers)
    for _ in range(args.warmup):
        step()
    barrier()

    # -- value: K steps, device-timed, barrier + synchronize on both sides, max over ranks
    sampler = ClockSampler(local_rank)
    sampler.start()
    b2.capi.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush_ms = 0.0
    barrier()
    if flush is None:
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
        elapsed_ms = e0.elapsed_time(e1)
    else:  # small working set: flush L2 before every step, time each step on its own
        elapsed_ms = 0.0
        for _ in range(args.steps):
            flush.zero_()
            e0.record()
            step()
            e1.record()
            torch.cuda.synchronize(dev)
            elapsed_ms += e0.elapsed_time(e1)
        barrier()
    launches = b2.capi.launch_count()
    clocks = sampler.finish()
    elapsed_ms = _max_over_ranks(elapsed_ms, dev, dist, world)
    ms_per_step = elapsed_ms / args.steps
    value = world * ops_step / (ms_per_step * 1e-3) / 1e12

    # -- value_sustained: the same step back to back for >= 2 s (the driver's K may give a 20 ms burst region)
    peaks = load_peaks()
    sus_steps = max(args.steps, int(2200.0 / max(ms_per_step, 1e-3)) + 1) if flush is None else 0
    value_sustained = None
    if sus_steps:
        barrier()
        se0, se1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        se0.record()
        for _ in range(sus_steps):
            step()
        se1.record()
        barrier()
        sus_ms = _max_over_ranks(se0.elapsed_time(se1), dev, dist, world) / sus_steps
        value_sustained = {"value": round(world * ops_step / (sus_ms * 1e-3) / 1e12, 1), "unit": "TOPS",
                           "ms_per_step": round(sus_ms, 4), "steps": sus_steps,
                           "frac_of_datasheet_int8_4500_tops": round(ops_step / (sus_ms * 1e-3) / 1e12 / DATASHEET_INT8_TOPS, 4)}

    # -- roofline: the GEMM kernel alone (quantized activations prepared outside the timed launches), two timing modes,
    #    each against the peak that matches it: (a) back-to-back launches for >= 2 s vs the SUSTAINED peak,
    #    (b) isolated launches (events + synchronize around each) vs the BURST peak
    qs = [b2.per_token_quant_int8(x) for x, _, _, _ in probs]
    gemm_ms = 0.0
    n_gemm = 0
    ge0, ge1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(3, min(args.steps, 50))
    for it in range(3 + reps):
        for (x, W, Bs, b), (q, s) in zip(probs, qs):
            if flush is not None:
                flush.zero_()
            ge0.record()
            b2.int8_scaled_mm(q, W.t(), s, Bs.view(-1, 1), torch.bfloat16, b)
            ge1.record()
            torch.cuda.synchronize(dev)
            if it >= 3:
                gemm_ms += ge0.elapsed_time(ge1)
                n_gemm += 1
    gemm_cfg = b2.capi.last_gemm_config()
    burst_tops = reps * ops_step / (gemm_ms * 1e-3) / 1e12
    burst_us = gemm_ms * 1e3 / n_gemm
    mma_peak = mma_only_peak()
    if flush is None:
        rounds = int(2200.0 / max(gemm_ms / reps, 1e-3)) + 1
        torch.cuda.synchronize(dev)
        ge0.record()
        for _ in range(rounds):
            for (x, W, Bs, b), (q, s) in zip(probs, qs):
                b2.int8_scaled_mm(q, W.t(), s, Bs.view(-1, 1), torch.bfloat16, b)
        ge1.record()
        torch.cuda.synchronize(dev)
        sus_gemm_ms = ge0.elapsed_time(ge1)
        gemm_tops = rounds * ops_step / (sus_gemm_ms * 1e-3) / 1e12
        peak_tops, peak_name = 2.0 * peaks["bf16_sustained"], "bf16_tflops_sustained"
        timing = f"{rounds * len(probs)} launches back to back ({sus_gemm_ms / 1e3:.2f} s), CUDA events on the launch stream"
        avg_us = sus_gemm_ms * 1e3 / (rounds * len(probs))
        mma_frac = gemm_tops / mma_peak["int8_sustained"]
    else:  # L2-resident working set: only isolated, flushed launches are meaningful
        gemm_tops, peak_tops, peak_name = burst_tops, 2.0 * peaks["bf16_burst"], "bf16_tflops (burst)"
        timing = f"{n_gemm} isolated launches, L2 flushed before each"
        avg_us = burst_us
        mma_frac = gemm_tops / mma_peak["int8_burst"]
    roofline = {"bound": "tensor", "kernel": f"gemm_w8a8_kernel<kind::i8> ({gemm_cfg})",
                "achieved": round(gemm_tops, 1), "peak": round(peak_tops, 1), "unit": "TFLOP/s",
                "frac": round(gemm_tops / peak_tops, 4), "traffic": ncu_traffic(args.workload),
                "timing": timing, "avg_launch_us": round(avg_us, 2),
                "peak_source": f"2 x {peak_name} of MEASURED_PEAKS.json ({peaks['source']}): the cuBLAS bf16 GEMM rate of "
                               "this pool doubled for 8-bit operands, matched to the timing mode",
                "frac_of_mma_only_peak": round(mma_frac, 4), "mma_only_peak": mma_peak,
                "frac_of_datasheet_4500": round(gemm_tops / DATASHEET_INT8_TOPS, 4),
                "burst": {"achieved": round(burst_tops, 1), "peak": round(2.0 * peaks["bf16_burst"], 1),
                          "frac": round(burst_tops / (2.0 * peaks["bf16_burst"]), 4),
                          "frac_of_mma_only_peak": round(burst_tops / mma_peak["int8_burst"], 4),
                          "avg_launch_us": round(burst_us, 2), "launches_timed": n_gemm,
                          "timing": "isolated launches, events + synchronize around each"}}
    del qs

    # -- e2e: host buffers in, host buffers out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        host_layers = []
        for (x, W, Bs, b) in probs:
            hl = b2.HostW8A8Int8Linear(W, Bs, b, device=dev, chunk_rows=min(int(os.environ.get("B2_E2E_CHUNK", "2048")), M))
            xh = torch.empty(x.shape, dtype=torch.bfloat16).pin_memory()
            xh.copy_(x)
            oh = torch.empty(M, W.shape[0], dtype=torch.bfloat16).pin_memory()
            host_layers.append((hl, xh, oh))
        e2e_steps = max(3, min(args.steps, 20))
        for _ in range(2):
            for hl, xh, oh in host_layers:
                hl(xh, oh)
        barrier()
        t0 = time.perf_counter()
        ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ee0.record()
        for _ in range(e2e_steps):
            # the layers of a step are independent problems (as in the device-resident step above): enqueue all of
            # them, then collect -- uploads of one layer run under the downloads of another
            for hl, xh, oh in host_layers:
                hl.submit(xh, oh)
            for hl, _, _ in host_layers:
                hl.wait()
        ee1.record()
        barrier()
        e2e_ms = max(ee0.elapsed_time(ee1), 0.0)
        wall_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms = max(e2e_ms, 0.0)
        tt = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item()) / e2e_steps
        h2d = sum(xh.numel() * 2 for _, xh, _ in host_layers)
        d2h = sum(oh.numel() * 2 for _, _, oh in host_layers)
        # check the host result of the last step against the device-resident path (bitwise)
        ref_out = step()
        same = all(torch.equal(oh.view(torch.int16), o.cpu().view(torch.int16)) for (_, _, oh), o in zip(host_layers, ref_out))
        # what the box's host <-> device links give at best: every rank copies 256 MiB pinned buffers in both directions
        # at once, all ranks together (the e2e pipeline above is bound by exactly this on one GPU and by the shared host
        # memory / PCIe root on 8)
        hb = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
        hb2 = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
        db, db2 = torch.empty(256 << 20, dtype=torch.uint8, device=dev), torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        for it in range(2):
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            s_in.wait_event(c0)
            s_out.wait_event(c0)
            for _ in range(4):
                with torch.cuda.stream(s_in):
                    db.copy_(hb, non_blocking=True)
                with torch.cuda.stream(s_out):
                    hb2.copy_(db2, non_blocking=True)
            torch.cuda.current_stream(dev).wait_stream(s_in)
            torch.cuda.current_stream(dev).wait_stream(s_out)
            c1.record()
            torch.cuda.synchronize(dev)
            ceil_ms = c0.elapsed_time(c1)
        ceil_ms = _max_over_ranks(ceil_ms, dev, dist, world)
        ceil_gbs = 8 * (256 << 20) / (ceil_ms * 1e-3) / 1e9  # per rank, both directions together
        # the step's own traffic is lopsided (1.7x more bytes come back than go in): the floor for THIS step is the time
        # the same byte counts take as plain concurrent copies, nothing else on the GPU
        def _copies(stream, dst, src, nbytes):
            with torch.cuda.stream(stream):
                left = nbytes
                while left > 0:
                    n = min(left, 256 << 20)
                    dst[:n].copy_(src[:n], non_blocking=True)
                    left -= n
        for it in range(2):
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            s_in.wait_event(c0)
            s_out.wait_event(c0)
            _copies(s_in, db, hb, h2d)
            _copies(s_out, hb2, db2, d2h)
            torch.cuda.current_stream(dev).wait_stream(s_in)
            torch.cuda.current_stream(dev).wait_stream(s_out)
            c1.record()
            torch.cuda.synchronize(dev)
            floor_ms = c0.elapsed_time(c1)
        floor_ms = _max_over_ranks(floor_ms, dev, dist, world)
        del hb, hb2, db, db2
        e2e = {"value": round(world * ops_step / (e2e_ms * 1e-3) / 1e12, 2), "unit": "TOPS",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": round(e2e_ms, 3),
               "steps": e2e_steps, "host_wall_ms_per_step": round(wall_ms / e2e_steps, 3),
               "pcie_gbs": round((h2d + d2h) / (e2e_ms * 1e-3) / 1e9, 1), "matches_device_path_bitwise": bool(same),
               "link_ceiling_gbs_per_rank": round(ceil_gbs, 1),
               "link_ceiling_note": "concurrent pinned H2D + D2H copies of 256 MiB on every rank at once, slowest rank",
               "frac_of_link_ceiling": round((h2d + d2h) / (e2e_ms * 1e-3) / 1e9 / ceil_gbs, 3),
               "link_floor_ms_per_step": round(floor_ms, 3),
               "link_floor_note": "the step's own h2d and d2h byte counts as plain concurrent pinned copies, every rank at once",
               "frac_of_link_floor": round(floor_ms / e2e_ms, 3)}
        del host_layers

    # -- cpu baseline (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tops, dt, kind, cores, n, sample, rows = cpu_reference_run(layers, has_bias, CPU_SAMPLE_ROWS, 2, 5, min_seconds=8.0)
        cpu = {"value": round(tops, 4), "unit": "TOPS", "cores": cores, "kind": kind, "sample": sample,
               "ms_per_step": round(dt * 1e3, 3)}

    # -- multi-GPU legs (SURVEY 8e rows the DP headline cannot show) + parity block
    multi = {}
    parity_ok = True
    if world > 1 and not args.no_multi:
        try:
            from sglang_diffusion_b200.comm import SymmAllReduce

            comm = SymmAllReduce(max_bytes=256 << 20, device=dev, full_bytes=128 << 20)  # two sets of M = 8192 x N = 8192 bf16 partials + the gathered sum
            SymmAllReduce.set_timeout_ms(20000)  # a lost rank must not hang the box; checked below
        except Exception as e:  # noqa: BLE001
            comm = None
            multi["symm_comm_error"] = f"{type(e).__name__}: {e}"
        multi["strong"] = strong_scaling_leg(b2, dev, dist, world, rank, args.steps)
        tp_rows, tp_check = tp_leg(b2, dev, dist, world, rank, comm)
        multi["tp"] = {"workload": "Llama-3-70B W8A8-fp8 decoder-layer linears (qkv 8192x10240 col, o_proj 8192x8192 row, "
                                   "gate_up 8192x57344 col, down 28672x8192 row), TP = n_gpus", "rows": tp_rows}
        par = {"tp_row_parallel_vs_fp64_shard_products": tp_check,
               "m_sharded": m_sharded_parity(b2, dev, dist, world, rank)}
        if comm is not None:
            par["symm_allreduce"] = symm_allreduce_parity(dev, dist, world, rank, comm)
            par["symm_allreduce"]["wait_timed_out_at_barrier"] = comm.timed_out()
            par["symm_allreduce"]["ok"] = par["symm_allreduce"]["ok"] and par["symm_allreduce"]["wait_timed_out_at_barrier"] == 0
        parity_ok = all(v.get("ok", False) for v in par.values())
        par["ok"] = parity_ok
        multi["multi_gpu_parity"] = par

    if rank == 0:
        line = {"metric": "w8a8_int8_linear_tops", "value": round(value, 1), "unit": "TOPS", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
                "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int8",
                "data": "synthetic", "config": config, "tokens_per_s": round(world * M / (ms_per_step * 1e-3), 1),
                "frac_of_datasheet_int8_4500_tops": round(value / world / DATASHEET_INT8_TOPS, 4),
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "decode_roofline": decode,
                "value_sustained": value_sustained, "cpu_baseline": cpu}
        line.update(multi)
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    if not parity_ok:
        sys.exit(3)


if __name__ == "__main__":
    main()
