"""Parity + timing of b2_allreduce_symm against NCCL (torchrun, one rank per GPU).  Exit code 0 = all checks passed.

  * one-shot (algo 1): equals the fp32 sum of the bf16 partials taken in rank order, rounded once -- bit for bit, and
    identical on every rank;
  * two-shot (algo 2, NVSwitch multimem): fp32 accumulation inside the switch, order unspecified -- within 1 bf16 ulp of
    the exact sum and identical on every rank;
  * 200 back-to-back calls with changing data (flag reuse), graph replay, sizes from 16 B to 4 MB.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sglang_diffusion_b200 as b2  # noqa: E402,F401
from sglang_diffusion_b200.comm import SymmAllReduce  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
BIG = "--big" in sys.argv  # also time prefill-sized messages (needs a 128 MiB symmetric buffer)
comm = SymmAllReduce(max_bytes=(128 << 20) if BIG else (4 << 20), device=dev)
algos = [1] + ([2] if comm.multicast_ptr else [])
fails = 0


def exact_sum(part):
    """fp32 sum in rank order of every rank's bf16 partial (gathered with NCCL), rounded once to bf16."""
    parts = [torch.empty_like(part) for _ in range(world)]
    dist.all_gather(parts, part.clone())
    acc = parts[0].float()
    for p in parts[1:]:
        acc = acc + p.float()
    return acc, acc.to(part.dtype)


for dtype in (torch.bfloat16, torch.float16):
    for M, N in ((1, 8), (1, 4096), (16, 8192), (64, 8192), (100, 5120), (128, 8192), (256, 8192)):
        g = torch.Generator(device=dev).manual_seed(1000 * rank + M + N)
        for algo in algos:
            part = comm.partial(M, N, dtype)
            part.copy_(torch.randn(M, N, generator=g, device=dev).to(dtype))
            acc32, want = exact_sum(part)
            got = comm.all_reduce(part, algo)
            torch.cuda.synchronize()
            if algo == 1:
                ok = torch.equal(got.view(torch.int16), want.view(torch.int16))
            else:
                ulp = torch.finfo(dtype).eps * acc32.abs().clamp_min(1e-3)
                ok = bool(((got.float() - acc32).abs() <= ulp).all())
            ref0 = got.clone()
            dist.broadcast(ref0, 0)
            same = torch.equal(ref0.view(torch.int16), got.view(torch.int16))
            if not (ok and same):
                fails += 1
                print(f"[rank {rank}] FAIL dtype={dtype} M={M} N={N} algo={algo} ok={ok} same_on_all_ranks={same}", flush=True)

if comm.multicast_ptr:  # in place: the sums are left in the symmetric buffer itself
    part = comm.partial(64, 8192, torch.bfloat16)
    part.copy_(torch.randn(64, 8192, device=dev).bfloat16())
    acc32, _ = exact_sum(part)
    got = comm.all_reduce(part, 2, inplace=True)
    torch.cuda.synchronize()
    if got.data_ptr() != part.data_ptr() or not bool(((got.float() - acc32).abs() <= torch.finfo(torch.bfloat16).eps * acc32.abs().clamp_min(1e-3)).all()):
        fails += 1
        print(f"[rank {rank}] FAIL in-place two-shot", flush=True)

# flag reuse: many calls back to back, new data every time, checked at the end through a running checksum
for algo in algos:
    M, N = 64, 8192
    g = torch.Generator(device=dev).manual_seed(77 + rank)
    total = torch.zeros(M, N, device=dev)
    want_total = torch.zeros(M, N, device=dev)
    for it in range(200):
        part = comm.partial(M, N, torch.bfloat16)
        part.copy_(torch.randn(M, N, generator=g, device=dev).bfloat16())
        ref = part.float().clone()
        dist.all_reduce(ref)  # NCCL fp32 reference of the same partials
        total += comm.all_reduce(part, algo).float()
        want_total += ref
    torch.cuda.synchronize()
    err = ((total - want_total).abs().max() / want_total.abs().max()).item()
    if err > 5e-3:  # 200 sums, each rounded to bf16 once, against an fp32 NCCL reduction
        fails += 1
        print(f"[rank {rank}] FAIL 200-call chain algo={algo} rel err {err}", flush=True)

# timing, graph-replayed: quant-free, the all-reduce alone vs NCCL on the same message
def timed(fn, inner=20, reps=20):
    fn()
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(inner):
                fn()
    torch.cuda.current_stream().wait_stream(st)
    gr.replay()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (reps * inner)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3


for M, N in ((1, 4096), (16, 8192), (64, 8192), (256, 8192)) + (((2048, 8192), (8192, 8192)) if BIG else ()):
    part = comm.partial(M, N, torch.bfloat16)
    part.normal_()
    nccl_buf = part.clone()
    row = {"world": world, "bytes": M * N * 2, "us_nccl": round(timed(lambda: dist.all_reduce(nccl_buf)), 2)}
    for algo in (algos if M <= 256 else [a for a in algos if a == 2]):
        row["us_one_shot_p2p" if algo == 1 else "us_two_shot_multimem"] = round(timed(lambda: comm.all_reduce(part, algo)), 2)
    if comm.multicast_ptr:
        row["us_two_shot_multimem_inplace"] = round(timed(lambda: comm.all_reduce(part, 2, inplace=True)), 2)
    if rank == 0:
        print(json.dumps(row), flush=True)

f = torch.tensor([fails], device=dev)
dist.all_reduce(f)
if rank == 0:
    print("SYMM_ALLREDUCE_CHECK", "PASS" if int(f.item()) == 0 else f"FAIL ({int(f.item())})", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(f.item()) == 0 else 1)
