"""Stress / diagnosis of the scatter epilogue: after b2_gemm_s8_scatter_symm, a device synchronize and a group barrier,
every slot of this rank's gather buffer must equal the owning rows of the producing rank's plain GEMM output.  Reports
which (slot, row, column) differ.  torchrun, one rank per GPU."""
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
comm = SymmAllReduce(max_bytes=72 << 20, device=dev)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
fails = 0
shapes = [(1024 + 32 * world, 1032, 384), (4096, 4096, 512), (1024 + 32 * world, 1032, 384), (1280, 1088, 384),
          (1280, 520, 256), (2048, 1032, 256)]
for it in range(iters):
    M, N, K = shapes[it % len(shapes)]
    if M % (32 * world):
        continue
    g = torch.Generator(device=dev).manual_seed(it * 977 + rank)
    x = (torch.randn(M, K, generator=g, device=dev) / 10).bfloat16()
    W = torch.randint(-127, 128, (N, K), dtype=torch.int8, generator=g, device=dev)
    q, s = ops.per_token_quant_int8(x)
    Bs = torch.rand(N, generator=g, device=dev) * 1e-2
    part = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    ops.scaled_mm_out(q, W.t(), s.view(-1), Bs, part, None)
    parts = [torch.empty_like(part) for _ in range(world)]
    dist.all_gather(parts, part)
    torch.cuda.synchronize()
    counter = int(comm.handle.get_signal_pad(rank, (1546,), torch.uint32)[1544].item())
    comm.scatter_gemm(q, W.t(), s.view(-1), Bs, torch.bfloat16, None)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    rows = M // world
    base = (counter & 1) * comm.set_stride
    slots = comm.buf[base:base + world * rows * N * 2].view(torch.bfloat16).view(world, rows, N).clone()
    for r in range(world):
        want = parts[r][rank * rows:(rank + 1) * rows]
        bad = (slots[r].view(torch.int16) != want.view(torch.int16)).nonzero()
        if bad.numel():
            fails += 1
            rr, cc = bad[:, 0], bad[:, 1]
            print(f"[rank {rank}] it={it} M={M} N={N} set={counter & 1} slot(from rank)={r}: {bad.shape[0]} wrong, rows "
                  f"{sorted(set(rr.tolist()))[:12]} cols {sorted(set(cc.tolist()))[:24]}", flush=True)
    out = comm.reduce_scattered(M, N, torch.bfloat16)  # advances the counter
    torch.cuda.synchronize()
t = torch.tensor([fails], device=dev)
dist.all_reduce(t)
if rank == 0:
    print("SYMM_SCATTER_STRESS PASS" if int(t.item()) == 0 else f"SYMM_SCATTER_STRESS FAIL ({int(t.item())})", flush=True)
dist.destroy_process_group()
