#!/usr/bin/env python
"""What bounds the end-to-end leg at N GPUs: pinned host -> device copies from N processes at once.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29517 tools/h2d_probe.py [--mb 255]

MEASUREMENT PROBE (not product code).  Every rank copies the same 255 MB (bench.py's C2 inputs) from pinned host memory
to its GPU in a loop, all ranks started together by a barrier; per-rank and aggregate GB/s.  Then the same with the
process pinned (sched_setaffinity) to the CPUs that sysfs lists as local to its GPU and the buffer allocated AFTER
pinning (first-touch places the pages on that NUMA node).  Rank 0 prints one JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist


def local_cpus(dev: int) -> list[int]:
    try:
        bus = torch.cuda.get_device_properties(dev).pci_bus_id
        dom = torch.cuda.get_device_properties(dev).pci_domain_id
        devid = torch.cuda.get_device_properties(dev).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{devid:02x}.0/local_cpulist"
        txt = open(path).read().strip()
    except Exception:
        return []
    cpus: list[int] = []
    for part in txt.split(","):
        if "-" in part:
            a, b = part.split("-")
            cpus += list(range(int(a), int(b) + 1))
        elif part:
            cpus.append(int(part))
    return cpus


def measure(nbytes: int, iters: int, world: int) -> float:
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.fill_(65)
    devb = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        devb.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        devb.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return nbytes * iters / (e0.elapsed_time(e1) * 1e-3) / 1e9


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=255)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = a.mb << 20

    def gather(v: float) -> list[float]:
        if world == 1:
            return [v]
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(x.item()) for x in out]

    plain = gather(measure(nbytes, a.iters, world))
    cpus = local_cpus(local)
    pinned = None
    if cpus:
        try:
            os.sched_setaffinity(0, set(cpus))
            time.sleep(0.01)
            pinned = gather(measure(nbytes, a.iters, world))
        except OSError:
            pinned = None
    else:
        gather(0.0) if world > 1 else None
    if rank == 0:
        topo = ""
        try:
            topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        except Exception:
            pass
        numa = [ln for ln in topo.splitlines() if ln.startswith("GPU")]
        print(json.dumps({"probe": "pinned_h2d", "n_procs": world, "mb_per_copy": a.mb, "iters": a.iters,
                          "gb_per_s_per_rank": [round(x, 2) for x in plain], "gb_per_s_aggregate": round(sum(plain), 1),
                          "affinity_local_cpus_rank0": cpus[:4] + (["..."] if len(cpus) > 4 else []),
                          "gb_per_s_per_rank_affinity": [round(x, 2) for x in pinned] if pinned else None,
                          "gb_per_s_aggregate_affinity": round(sum(pinned), 1) if pinned else None,
                          "host_cores": os.cpu_count(), "topo_gpu_rows": numa[:8]}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
