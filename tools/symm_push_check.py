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
