"""Probe: is torch symmetric memory (peer pointers + multicast) usable on this box?  torchrun, one rank per GPU."""
import os, sys
import torch, torch.distributed as dist
rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm_mem
    t = symm_mem.empty(1 << 20, dtype=torch.bfloat16, device=dev)
    h = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "rendezvous ok", type(h).__name__, "world", h.world_size, "rank", h.rank, flush=True)
    for a in ("buffer_ptrs", "signal_pad_ptrs", "multicast_ptr", "buffer_size", "signal_pad_size", "buffer_ptrs_dev", "signal_pad_ptrs_dev"):
        try:
            v = getattr(h, a)
            print(rank, a, (v if not isinstance(v, (list, tuple)) else [hex(x) for x in v]), flush=True)
        except Exception as e:
            print(rank, a, "ERR", e, flush=True)
    t.fill_(rank + 1); torch.cuda.synchronize(); dist.barrier()
    peer = h.get_buffer((rank + 1) % h.world_size, (16,), torch.bfloat16)
    print(rank, "peer value", peer[0].item(), flush=True)
    print(rank, "can_access_peer", torch.cuda.can_device_access_peer(local, (local + 1) % torch.cuda.device_count()), flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
    print(rank, "SYMM FAILED", type(e).__name__, e, flush=True)
dist.barrier(); dist.destroy_process_group()
