"""Parity + timing of the fused reduce-scatter of the prefill row-parallel layer (b2_gemm_*_scatter_symm: the
persistent GEMM's epilogue stores every rank's rows into that rank's gather buffer over NVLink; then
b2_allreduce_symm_gathered sums the slots) against the separate path (GEMM into the symmetric buffer + one-shot
b2_allreduce_symm, algo 1, rows of this rank): the same rank-ordered fp32 sum, so equal bit for bit.  Also mixes
decode-sized push calls in between (the two forms share the set counter) and replays an odd number of calls in a graph.
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
TIMING = "--time" in sys.argv
comm_rs = SymmAllReduce(max_bytes=(160 << 20) if TIMING else (72 << 20), device=dev,
                        full_bytes=(80 << 20) if TIMING else (36 << 20))
comm_ref = SymmAllReduce(max_bytes=(64 << 20) if TIMING else (32 << 20), device=dev)
fails = 0


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


LAST_FULL = [None]


def reference(q, s, W, Bs, bias, M, N, dtype=torch.bfloat16):
    part = ops.scaled_mm_out(q, W.t(), s, Bs, comm_ref.partial(M, N, dtype), bias)
    full = comm_ref.all_reduce(part, 1)
    LAST_FULL[0] = full
    r = M // world
    return full[rank * r:(rank + 1) * r]


shapes = ((32 * world, 512, 256), (256 * world, 512, 1024), (1024, 4096, 1792), (2048, 5120, 640), (4096, 4096, 512),
          (1024 + 32 * world, 1032, 384))
for kind in ("int8", "fp8"):
    for M, N, K in shapes:
        if M % (32 * world):
            continue
        for it in range(3):  # both slot sets and the wrap-around
            q, s, W, Bs, bias = problem(kind, M, N, K, M + N + it)
            want = reference(q, s, W, Bs, bias, M, N)
            assert comm_rs.fits_scatter(M, N, torch.bfloat16)
            comm_rs.scatter_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            got = comm_rs.reduce_scattered(M, N, torch.bfloat16)
            torch.cuda.synchronize()
            ok = torch.equal(got.view(torch.int16), want.view(torch.int16))
            if not ok:
                fails += 1
                bad = (got.view(torch.int16) != want.view(torch.int16)).sum().item()
                print(f"[rank {rank}] FAIL {kind} M={M} N={N} K={K} call={it} mismatches={bad}/{got.numel()} "
                      f"cfg={capi.last_gemm_config()}", flush=True)
            if comm_rs.fits_scatter_all_reduce(M, N, torch.bfloat16):  # all-reduce form: + multicast all-gather
                comm_rs.scatter_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
                full = comm_rs.all_reduce_scattered(M, N, torch.bfloat16, inplace=(it == 2))
                torch.cuda.synchronize()
                if not torch.equal(full.view(torch.int16), LAST_FULL[0].view(torch.int16)):
                    fails += 1
                    print(f"[rank {rank}] FAIL all-reduce form {kind} M={M} N={N} K={K} call={it}", flush=True)
                dist.barrier()  # the in-place view is read above: keep the peers from overwriting it early
            if it == 1 and comm_rs.multicast_ptr:  # a decode-sized push call in between: different size, same counter
                qd, sd, Wd, Bsd, biasd = problem("int8", 16, 2048, 1024, 7 + it)
                wd = ops.scaled_mm_out(qd, Wd.t(), sd, Bsd, comm_ref.partial(16, 2048, torch.bfloat16), biasd)
                wd = comm_ref.all_reduce(wd, 1)
                comm_rs.push_gemm(qd, Wd.t(), sd, Bsd, torch.bfloat16, biasd)
                gd = comm_rs.reduce_gathered(16, 2048, torch.bfloat16)
                torch.cuda.synchronize()
                if not torch.equal(gd.view(torch.int16), wd.view(torch.int16)):
                    fails += 1
                    print(f"[rank {rank}] FAIL interleaved push call after {kind} M={M}", flush=True)

# reduce-scatter + residual add + RMSNorm + quantizer + all-gather of the codes, one launch behind the scatter GEMM:
# against the separate steps (rows of the one-shot sum -> ops.rmsnorm_quant on this rank's rows -> NCCL all-gather)
for mode in ("int8", "fp8"):
    for M, N, K in ((32 * world, 512, 256), (1024, 4096, 512), (2048, 5120, 640), (1024 + 32 * world, 1040, 384)):
        if M % (32 * world) or not comm_rs.fits_scatter_rmsnorm_allgather(M, N, torch.bfloat16):
            continue
        for it in range(2):
            q, s, W, Bs, bias = problem("int8", M, N, K, 3 * M + N + it)
            rows = M // world
            g = torch.Generator(device=dev).manual_seed(41 + it)  # same on every rank
            wnorm = (torch.rand(N, generator=g, device=dev) + 0.5).bfloat16()
            res_full = torch.randn(M, N, generator=g, device=dev).bfloat16()
            res_a = res_full[rank * rows:(rank + 1) * rows].clone() if it == 0 else None
            res_b = None if res_a is None else res_a.clone()
            want_rows = reference(q, s, W, Bs, bias, M, N).clone()
            qa, sa = ops.rmsnorm_quant(want_rows, wnorm, 1e-6, residual=res_a, mode=mode)
            qa_all = torch.empty(M, N, dtype=torch.uint8, device=dev)
            sa_all = torch.empty(M, 1, dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(qa_all, qa.view(torch.uint8).contiguous())
            dist.all_gather_into_tensor(sa_all, sa.contiguous())
            comm_rs.scatter_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            qb, sb = comm_rs.reduce_scattered_rmsnorm_quant_allgather(M, N, torch.bfloat16, wnorm, 1e-6,
                                                                      residual_local=res_b, mode=mode)
            torch.cuda.synchronize()
            ok = (torch.equal(qa_all, qb.view(torch.uint8)) and torch.equal(sa_all, sb)
                  and (res_a is None or torch.equal(res_a.view(torch.int16), res_b.view(torch.int16))))
            if not ok:
                fails += 1
                nbad = (qa_all != qb.view(torch.uint8)).sum().item()
                print(f"[rank {rank}] FAIL rs+rmsnorm+quant+allgather {mode} M={M} N={N} it={it} bad codes {nbad} "
                      f"bad scales {(sa_all != sb).sum().item()}", flush=True)
            dist.barrier()  # the views live in the shared buffer: keep the peers' next call behind this rank's reads

# f16 output once
M, N, K = 64 * world, 1024, 512
g = torch.Generator(device=dev).manual_seed(5 + rank)
x = (torch.randn(M, K, generator=g, device=dev) / 10).half()
W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
q, s = ops.per_token_quant_int8(x)
Bs = torch.rand(N, generator=g, device=dev) * 1e-2
want = reference(q, s.view(-1), W, Bs, None, M, N, torch.float16)
comm_rs.scatter_gemm(q, W.t(), s.view(-1), Bs, torch.float16, None)
got = comm_rs.reduce_scattered(M, N, torch.float16)
torch.cuda.synchronize()
if not torch.equal(got.view(torch.int16), want.view(torch.int16)):
    fails += 1
    print(f"[rank {rank}] FAIL f16 output", flush=True)

# graph replay, odd number of calls per graph
M, N, K = 1024, 4096, 1024
q, s, W, Bs, bias = problem("int8", M, N, K, 999)
want = reference(q, s, W, Bs, bias, M, N).clone()
torch.cuda.synchronize()
st = torch.cuda.Stream()
st.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(st):
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=st):
        for _ in range(3):
            comm_rs.scatter_gemm(q, W.t(), s, Bs, torch.bfloat16, bias)
            out = comm_rs.reduce_scattered(M, N, torch.bfloat16)
torch.cuda.current_stream().wait_stream(st)
for rep in range(5):
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    if not torch.equal(out.view(torch.int16), want.view(torch.int16)):
        fails += 1
        print(f"[rank {rank}] FAIL graph replay {rep}", flush=True)
if comm_rs.timed_out() or comm_ref.timed_out():
    fails += 1
    print(f"[rank {rank}] FAIL a barrier wait timed out", flush=True)

if TIMING:
    def timed(fn, iters=20):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for M, N, Kfull in ((8192, 4096, 14336), (8192, 4096, 4096), (4096, 5120, 13824), (2048, 4096, 14336),
                        (1024, 8192, 8192), (512, 4096, 14336), (256, 8192, 8192)):
        K = Kfull // world // 128 * 128
        q, s, W, Bs, bias = problem("int8", M, N, K, 1)
        Wt = W.t()
        plain = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        rs_out = torch.empty(M // world, N, dtype=torch.bfloat16, device=dev)

        def gemm_only():
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)

        def nccl_ar():
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)
            dist.all_reduce(plain)

        def nccl_rs():
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)
            dist.reduce_scatter_tensor(rs_out, plain)

        def symm_ar():
            part = ops.scaled_mm_out(q, Wt, s, Bs, comm_ref.partial(M, N, torch.bfloat16), bias)
            comm_ref.all_reduce(part)

        def fused_rs():
            comm_rs.scatter_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            comm_rs.reduce_scattered(M, N, torch.bfloat16)

        def scatter_only():  # the same set every time (the counter only advances in the sum): timing only
            comm_rs.scatter_gemm(q, Wt, s, Bs, torch.bfloat16, bias)

        def fused_ar():
            comm_rs.scatter_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            comm_rs.all_reduce_scattered(M, N, torch.bfloat16)

        def fused_ar_inplace():
            comm_rs.scatter_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            comm_rs.all_reduce_scattered(M, N, torch.bfloat16, inplace=True)

        extra = (("fused_scatter_gemm+sum+allgather", fused_ar), ("same_in_place", fused_ar_inplace)) \
            if comm_rs.fits_scatter_all_reduce(M, N, torch.bfloat16) else ()
        gq = torch.Generator(device=dev).manual_seed(3)
        wnorm = (torch.rand(N, generator=gq, device=dev) + 0.5).bfloat16()
        res_full = torch.randn(M, N, generator=gq, device=dev).bfloat16()
        res_loc = res_full[rank * (M // world):(rank + 1) * (M // world)].clone()
        q_all = torch.empty(M, N, dtype=torch.int8, device=dev)
        s_all = torch.empty(M, 1, dtype=torch.float32, device=dev)

        def nccl_ar_norm():  # all-reduce, then norm + quantizer of all M rows on every rank
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)
            dist.all_reduce(plain)
            ops.rmsnorm_quant(plain, wnorm, 1e-6, residual=res_full, mode="int8")

        def nccl_rs_norm_ag():  # sequence-parallel norm with NCCL collectives
            ops.scaled_mm_out(q, Wt, s, Bs, plain, bias)
            dist.reduce_scatter_tensor(rs_out, plain)
            ql, sl = ops.rmsnorm_quant(rs_out, wnorm, 1e-6, residual=res_loc, mode="int8")
            dist.all_gather_into_tensor(q_all, ql)
            dist.all_gather_into_tensor(s_all, sl)

        def fused_rs_norm_ag():
            comm_rs.scatter_gemm(q, Wt, s, Bs, torch.bfloat16, bias)
            comm_rs.reduce_scattered_rmsnorm_quant_allgather(M, N, torch.bfloat16, wnorm, 1e-6, residual_local=res_loc, mode="int8")

        if comm_rs.fits_scatter_rmsnorm_allgather(M, N, torch.bfloat16):
            extra = extra + (("gemm+nccl_all_reduce+rmsnorm_quant", nccl_ar_norm),
                             ("gemm+nccl_rs+rmsnorm_quant+nccl_ag", nccl_rs_norm_ag),
                             ("scatter_gemm+fused_sum_rmsnorm_quant_allgather", fused_rs_norm_ag))
        res = {name: timed(fn) for name, fn in extra + (("gemm_only", gemm_only), ("scatter_gemm_only", scatter_only), ("gemm+nccl_all_reduce", nccl_ar),
                                                ("gemm+nccl_reduce_scatter", nccl_rs), ("gemm+symm_all_reduce", symm_ar),
                                                ("fused_scatter_gemm+sum", fused_rs))}
        if rank == 0:
            nv = (world - 1) / world * M * N * 2
            print(f"TIMING world={world} M={M} N={N} K/rank={K}: " + "  ".join(f"{k} {v:.1f} us" for k, v in res.items())
                  + f"  | NVLink bytes out per rank {nv / 1e6:.1f} MB -> {nv / res['fused_scatter_gemm+sum'] / 1e3:.0f} GB/s over the fused call",
                  flush=True)

t = torch.tensor([fails], device=dev)
dist.all_reduce(t)
if rank == 0:
    print("SYMM_SCATTER_CHECK PASS" if int(t.item()) == 0 else f"SYMM_SCATTER_CHECK FAIL ({int(t.item())} failures)", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
