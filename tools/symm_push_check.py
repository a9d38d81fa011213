"""Parity of the GEMM-epilogue-fused all-reduce (b2_gemm_*_push_symm + b2_allreduce_symm_gathered) against the
separate path (GEMM into the symmetric buffer + one-shot b2_allreduce_symm, algo 1): same rank-ordered fp32 sum, so
the outputs must be equal bit for bit, on every rank, across back-to-back calls (the two slot sets alternate) and under
CUDA-graph replay with an ODD number of calls per graph (the set parity then differs between replays).
torchrun, one rank per GPU.  Exit code 0 = all checks passed."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2  # noqa: E402
from sglang_diffusion_b200 import capi, ops  # noqa: E402
from sglang_diffusion_b200.comm import SymmAllReduce  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
SymmAllReduce.set_timeout_ms(20000)
comm_push = SymmAllReduce(max_bytes=32 << 20, device=dev)   # pushed slots
comm_ref = SymmAllReduce(max_bytes=8 << 20, device=dev)     # separate-kernel reference path
fails = 0
if not comm_push.multicast_ptr:
    if rank == 0:
        print("SYMM_PUSH_CHECK SKIP (no multicast mapping on this box)")
    dist.destroy_process_group()
    sys.exit(0)


def problem(kind, M, N, K, seed):
    g = torch.Generator(device=dev).manual_seed(seed * 131 + rank)
    x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
    if kind == "int8":
        W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
        q, s = ops.per_token_quant_int8(x)
    else:
        W = (torch.randn(N, K, generator=g, device=dev)).to(torch.float8_e4m3fn)
        q, s = ops.sglang_per_token_quant_fp8(x)
    Bs = torch.rand(N, generator=g, device=dev) * 1e-2
    bias = torch.randn(N, generator=g, device=dev).bfloat16() if rank == 0 else None
    return q, s.view(-1), W, Bs, bias


def reference(q, s, W, Bs, bias, M, N):
    part = ops.scaled_mm_out(q, W.t(), s, Bs, comm_ref.partial(M, N, torch.bfloat16), bias)
    return comm_ref.all_reduce(part, 1)


for kind in ("int8", "fp8"):
    for M, N, K in ((1, 512, 1024), (16, 8192, 1024), (64, 8192, 3584), (64, 5120, 2048), (128, 4096, 4096)):
        for it in range(3):  # three calls per shape: both slot sets and the wrap-around
            q, s, W, Bs, bias = problem(kind, M, N, K, M + N + it)
            want = reference(q, s, W, Bs, bias, M, N)
            comm_push.push_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            got = comm_push.reduce_gathered(M, N, torch.bfloat16)
            torch.cuda.synchronize()
            ok = torch.equal(got.view(torch.int16), want.view(torch.int16))
            ref0 = got.clone()
            dist.broadcast(ref0, 0)
            same = torch.equal(ref0.view(torch.int16), got.view(torch.int16))
            if not (ok and same):
                fails += 1
                print(f"[rank {rank}] FAIL {kind} M={M} N={N} K={K} call={it} equal_to_one_shot={ok} same_on_all_ranks={same} "
                      f"cfg={capi.last_gemm_config()}", flush=True)

# all-reduce + fused_add_rmsnorm + quantizer in one launch == gathered sum, then ops.rmsnorm_quant, bit for bit
for mode in ("int8", "fp8", "fp8_group128"):
    for M, N, K in ((1, 512, 1024), (16, 8192, 1024), (64, 8192, 3584), (64, 5120, 2048), (100, 4096, 512)):
        if not comm_push.fits_reduce_rmsnorm(M, N, torch.bfloat16):
            continue
        for it in range(2):
            q, s, W, Bs, bias = problem("int8", M, N, K, 31 * M + N + it)
            g = torch.Generator(device=dev).manual_seed(5 + it)  # same on every rank
            wnorm = (torch.rand(N, generator=g, device=dev) + 0.5).bfloat16()
            res0 = torch.randn(M, N, generator=g, device=dev).bfloat16() if it == 0 else None
            # reference: the two separate launches
            comm_push.push_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            summed = comm_push.reduce_gathered(M, N, torch.bfloat16)
            res_a = None if res0 is None else res0.clone()
            qa, sa, ya = ops.rmsnorm_quant(summed, wnorm, 1e-6, residual=res_a, mode=mode, return_normed=True)
            # fused
            comm_push.push_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            res_b = None if res0 is None else res0.clone()
            qb, sb, yb = comm_push.reduce_gathered_rmsnorm_quant(M, N, torch.bfloat16, wnorm, 1e-6, residual=res_b,
                                                                 mode=mode, return_normed=True)
            torch.cuda.synchronize()
            ok = (torch.equal(qa.view(torch.uint8), qb.view(torch.uint8)) and torch.equal(sa, sb)
                  and torch.equal(ya.view(torch.int16), yb.view(torch.int16))
                  and (res0 is None or torch.equal(res_a.view(torch.int16), res_b.view(torch.int16))))
            if not ok:
                fails += 1
                print(f"[rank {rank}] FAIL fused rmsnorm {mode} M={M} N={N} K={K} it={it}", flush=True)

if "--time" in sys.argv:
    def timed(fn, iters=200):
        for _ in range(5):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        st_ = torch.cuda.Stream()
        st_.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st_):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st_):
                for _ in range(10):
                    fn()
        torch.cuda.current_stream().wait_stream(st_)
        gr.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters // 10):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / (iters // 10 * 10) * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for M, N, K in ((16, 8192, 1024), (64, 8192, 1024), (16, 4096, 1792), (64, 8192, 3584)):
        q, s, W, Bs, bias = problem("int8", M, N, K, 77)
        Wt = W.t()
        g = torch.Generator(device=dev).manual_seed(9)
        wnorm = (torch.rand(N, generator=g, device=dev) + 0.5).bfloat16()
        res = torch.randn(M, N, generator=g, device=dev).bfloat16()
        plain = torch.empty(M, N, dtype=torch.bfloat16, device=dev)

        def nccl_path():
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)
            dist.all_reduce(plain)
            ops.rmsnorm_quant(plain, wnorm, 1e-6, residual=res, mode="int8")

        def two_shot_path():
            part = ops.scaled_mm_out(q, Wt, s, Bs, comm_ref.partial(M, N, torch.bfloat16), bias)
            ops.rmsnorm_quant(comm_ref.all_reduce(part), wnorm, 1e-6, residual=res, mode="int8")

        def push_path():
            comm_push.push_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            ops.rmsnorm_quant(comm_push.reduce_gathered(M, N, torch.bfloat16), wnorm, 1e-6, residual=res, mode="int8")

        def fused_path():
            comm_push.push_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            comm_push.reduce_gathered_rmsnorm_quant(M, N, torch.bfloat16, wnorm, 1e-6, residual=res, mode="int8")

        with capi.static_weights():
            r = {n: timed(f) for n, f in (("gemm+nccl+rmsnorm_quant", nccl_path), ("gemm+two_shot+rmsnorm_quant", two_shot_path),
                                          ("push_gemm+sum+rmsnorm_quant", push_path), ("push_gemm+fused_sum_rmsnorm_quant", fused_path))}
        if rank == 0:
            print(f"TIMING world={world} M={M} N={N} K/rank={K}: " + "  ".join(f"{k} {v:.1f} us" for k, v in r.items()), flush=True)

# graph replay with an odd number of calls per graph
# (int8: exact int32 accumulation, so the comparison does not depend on which kernel plan either path picks)
M, N, K = 64, 8192, 1024
q, s, W, Bs, bias = problem("int8", M, N, K, 999)
want = reference(q, s, W, Bs, bias, M, N)
torch.cuda.synchronize()
st = torch.cuda.Stream()
st.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(st):
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=st):
        for _ in range(3):
            comm_push.push_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            out = comm_push.reduce_gathered(M, N, torch.bfloat16)
torch.cuda.current_stream().wait_stream(st)
for rep in range(5):
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    if not torch.equal(out.view(torch.int16), want.view(torch.int16)):
        fails += 1
        print(f"[rank {rank}] FAIL graph replay {rep}", flush=True)
if comm_push.timed_out() or comm_ref.timed_out():
    fails += 1
    print(f"[rank {rank}] FAIL a barrier wait timed out", flush=True)

t = torch.tensor([fails], device=dev)
dist.all_reduce(t)
if rank == 0:
    print("SYMM_PUSH_CHECK PASS" if int(t.item()) == 0 else f"SYMM_PUSH_CHECK FAIL ({int(t.item())} failures)", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
