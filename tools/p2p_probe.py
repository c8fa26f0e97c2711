#!/usr/bin/env python
"""Diagnostic (not a bench): what does rank-to-rank transport look like on this box?
torchrun --nproc-per-node N tools/p2p_probe.py  -- NCCL all_to_all_single rate, symmetric-memory
rendezvous (peer pointers usable from our own kernels), peer copy rate."""
import os, sys, time, json
import torch, torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
out = {"rank": rank, "world": world}

def timed(fn, n=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for mb in (8, 64, 340):
    n = mb * (1 << 20) // 8 // world * world
    send = torch.empty(n, dtype=torch.int64, device=dev).fill_(rank)
    recv = torch.empty_like(send)
    ms = timed(lambda: dist.all_to_all_single(recv, send))
    out[f"a2a_{mb}MB_ms"] = round(ms, 3)
    out[f"a2a_{mb}MB_GBs_out_per_rank"] = round(n * 8 * (world - 1) / world / ms / 1e6, 1)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(64 << 20, dtype=torch.uint8, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    out["symm_ok"] = True
    out["symm_ptrs"] = [hex(p) for p in hdl.buffer_ptrs][:4]
    peer = (rank + 1) % world
    pt = hdl.get_buffer(peer, (64 << 20,), torch.uint8)
    src = torch.empty(64 << 20, dtype=torch.uint8, device=dev).fill_(rank + 1)
    ms = timed(lambda: pt.copy_(src))
    out["symm_peer_copy_64MB_ms"] = round(ms, 3)
    out["symm_peer_copy_GBs"] = round((64 << 20) / ms / 1e6, 1)
    hdl.barrier()
    out["peer_wrote"] = int(t[0].item())
except Exception as e:  # noqa: BLE001
    out["symm_ok"] = False
    out["symm_err"] = repr(e)[:300]
try:
    out["can_access_peer"] = [torch.cuda.can_device_access_peer(lr, j) for j in range(torch.cuda.device_count()) if j != lr]
except Exception as e:  # noqa: BLE001
    out["can_access_peer"] = repr(e)[:100]
print(json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
