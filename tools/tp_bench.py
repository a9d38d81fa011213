"""BASELINE.json configs[4]: Llama-3-70B W8A8-fp8 tensor-parallel shards, decode M=64 and prefill M=8192.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/tp_bench.py [--int8] [--check]

One process per GPU (TP = N).  Column-parallel layers (QKV 8192 -> 10240, gate_up 8192 -> 57344) split W on N and
need no collective; row-parallel layers (o_proj 8192 -> 8192, down 28672 -> 8192) split W on K, quantize the local
K shard per token, and sum the partial outputs with ONE NCCL all-reduce on the compute stream
(python/sglang/srt/layers/linear.py:1467-1496 through sglang_diffusion_b200.linear.QuantLinear).
Every number is device time (CUDA events) per call, MAX over ranks; ops = 2*M*N*K of the WHOLE layer, so TOPS is
the job aggregate.  Prints one JSON line per (layer, M) on rank 0.
"""
from __future__ import annotations

import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sglang_diffusion_b200 as b2  # noqa: E402
from sglang_diffusion_b200.linear import QuantLinear, W8A8Fp8LinearMethod, W8A8Int8LinearMethod  # noqa: E402

LAYERS = [  # name, parallel, K (full), N (full)
    ("qkv_proj", "column", 8192, 10240),
    ("o_proj", "row", 8192, 8192),
    ("gate_up_proj", "column", 8192, 57344),
    ("down_proj", "row", 28672, 8192),
]


def timed(fn, iters, warm=5, inner=10):
    """Device time per call, max over ranks.  The calls are replayed from a CUDA graph (`inner` calls per graph,
    NCCL all-reduce included) so that decode shapes are not timed through the Python dispatcher; falls back to
    eager launches if the capture fails."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    run, per = fn, 1
    if "--eager" not in sys.argv:
        try:
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    for _ in range(inner):
                        fn()
            torch.cuda.current_stream().wait_stream(st)
            g.replay()
            torch.cuda.synchronize()
            run, per = g.replay, inner
        except Exception as e:  # noqa: BLE001
            print(f"[tp_bench] graph capture failed ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
            torch.cuda.synchronize()
    reps = max(1, iters // per)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (reps * per)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3


def main():
    int8 = "--int8" in sys.argv
    check = "--check" in sys.argv
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    # --as-tp N: use the shard shapes of TP = N on however many ranks are present (shape debugging on one GPU;
    # the all-reduce then only spans the ranks that exist)
    shard = world
    if "--as-tp" in sys.argv:
        shard = int(sys.argv[sys.argv.index("--as-tp") + 1])
    method = W8A8Int8LinearMethod() if int8 else W8A8Fp8LinearMethod()
    comm = None
    if "--symm" in sys.argv and world > 1:  # decode-sized row-parallel sums through b2_allreduce_symm (peer memory)
        from sglang_diffusion_b200.comm import SymmAllReduce
        # --prefill-symm: room for the scatter-epilogue all-reduce of the M = 8192 row-parallel layers as well
        big = "--prefill-symm" in sys.argv
        comm = SymmAllReduce(max_bytes=(272 << 20) if big else (4 << 20), device=dev, full_bytes=(136 << 20) if big else 0)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    for name, par, K, N in LAYERS:
        layer = QuantLinear(K, N, method, bias=False, parallel=par, tp_rank=rank, tp_size=shard, device=dev)
        layer.tp_size = world  # the collective spans the ranks that exist
        if comm is not None and par == "row":
            layer.use_symmetric_all_reduce(comm)
        n_loc, k_loc = (N // shard, K) if par == "column" else (N, K // shard)
        w = torch.randn(n_loc, k_loc, generator=g, device=dev) * 0.05
        ws = w.abs().amax(dim=1, keepdim=True) / (127.0 if int8 else 448.0)
        wq = torch.round(w / ws).clamp(-127, 127).to(torch.int8) if int8 else (w / ws).to(torch.float8_e4m3fn)
        layer.weight.data.copy_(wq)
        layer.weight_scale.data.copy_(ws)
        method.process_weights_after_loading(layer)
        del w
        for M in (64, 8192):
            x = (torch.randn(M, k_loc, generator=g, device=dev) / 10).bfloat16()
            out = layer(x)
            if check:
                part = (x.double() @ (wq.double() * ws.double()).t())
                if par == "row" and world > 1:
                    dist.all_reduce(part)
                err = ((out.double() - part).abs().max() / part.abs().max()).item()
                # dynamic 8-bit activations against the exact product of the same quantized weights: e4m3 keeps 3
                # significand bits (<= 6 % per element), int8 7 bits
                assert err < (3e-2 if int8 else 6e-2), (name, M, err)
            if check and par == "row" and world > 1 and M == 64:
                # row-parallel layer + all-reduce + fused_add_rmsnorm + next quantizer (two launches with --symm) against
                # the unfused composition: same sums up to the all-reduce's summation order (<= 1 bf16 ulp per element)
                gn = torch.Generator(device=dev).manual_seed(77)
                wn = (torch.rand(N, generator=gn, device=dev) + 0.5).bfloat16()
                res = torch.randn(M, N, generator=gn, device=dev).bfloat16()
                r1, r2 = res.clone(), res.clone()
                mode = "int8" if int8 else "fp8"
                q1, s1, y1 = layer.forward_allreduce_rmsnorm_quant(x, wn, 1e-6, residual=r1, mode=mode, return_normed=True)
                q2, s2, y2 = b2.ops.rmsnorm_quant(layer(x), wn, 1e-6, residual=r2, mode=mode, return_normed=True)
                torch.cuda.synchronize()
                assert (r1.float() - r2.float()).abs().max().item() <= 2.0 ** -6 * r2.float().abs().max().item(), name
                assert (y1.float() - y2.float()).abs().max().item() <= 2.0 ** -5 * y2.float().abs().max().item(), name
                d1, d2 = q1.float() * s1, q2.float() * s2
                assert (d1 - d2).abs().max().item() <= 0.08 * d2.abs().max().item(), name
            if check and par == "row" and world > 1 and M % world == 0:
                # sequence-parallel norm form against the all-reduce composition on this rank's rows, and the
                # all-gathered codes equal on every rank
                gn = torch.Generator(device=dev).manual_seed(78)
                wn = (torch.rand(N, generator=gn, device=dev) + 0.5).bfloat16()
                rows = M // world
                res = torch.randn(M, N, generator=gn, device=dev).bfloat16()[rank * rows:(rank + 1) * rows]
                r1, r2 = res.clone(), res.clone()
                mode = "int8" if int8 else "fp8"
                full = layer(x).clone()
                q1, s1 = layer.forward_reduce_scatter_rmsnorm_quant(x, wn, 1e-6, residual_local=r1, mode=mode)
                q1, s1 = q1.clone(), s1.clone()
                q2, s2 = b2.ops.rmsnorm_quant(full[rank * rows:(rank + 1) * rows].contiguous(), wn, 1e-6, residual=r2, mode=mode)
                torch.cuda.synchronize()
                assert (r1.float() - r2.float()).abs().max().item() <= 2.0 ** -6 * r2.float().abs().max().item(), name
                d1 = q1[rank * rows:(rank + 1) * rows].float() * s1[rank * rows:(rank + 1) * rows]
                d2 = q2.float() * s2
                assert (d1 - d2).abs().max().item() <= 0.08 * d2.abs().max().item(), name
                q0 = q1.view(torch.uint8).clone()
                dist.broadcast(q0, 0)
                assert torch.equal(q0, q1.view(torch.uint8)), name
                dist.barrier()
            iters = 200 if M == 64 else 30
            t_all = timed(lambda: layer(x), iters)
            t_mm = timed(lambda: method.apply(layer, x, None), iters)
            if rank == 0:
                ops = 2.0 * M * N * K
                print(json.dumps({
                    "layer": name, "parallel": par, "tp": world, "dtype": "int8" if int8 else "fp8_e4m3", "M": M,
                    "K_full": K, "N_full": N, "us_linear_incl_allreduce": round(t_all * 1e6, 1),
                    "us_quant_plus_gemm": round(t_mm * 1e6, 1),
                    "us_allreduce": round((t_all - t_mm) * 1e6, 1) if par == "row" else 0.0,
                    "allreduce_bytes": M * N * 2 if par == "row" and world > 1 else 0,
                    "tops_aggregate": round(ops / t_all / 1e12, 1),
                    "allreduce": (("b2_scatter_gemm+sum+allgather" if comm is not None and M * N * 2 >= ((8 << 20) if world < 8 else (32 << 20))
                                   and comm.fits_scatter_all_reduce(M, N, torch.bfloat16) else
                                   "b2_allreduce_symm" if comm is not None and comm.fits(M, N, torch.bfloat16)
                                   else "nccl") if par == "row" and world > 1 else None),
                    "gemm_config": b2.capi.last_gemm_config()}), flush=True)
        del layer
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
