"""D2H bandwidth per rank, alone and with all ranks copying at once (why does the per-step write-back of Vp
collapse at N = 8?).  torchrun --nproc-per-node N scripts/d2h_bw.py  ->  one JSON line on rank 0.
Each rank copies a 33.6 MB device buffer (C2's Vp) into pinned host memory 20 times; `alone` = ranks take turns, `together` = all
ranks start after a barrier.  Also reports the NUMA node of the pinned buffer's pages where /proc/self/numa_maps shows it."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N = 2099201 * 2
dev = torch.randn(N, dtype=torch.float64, device="cuda")
host = torch.empty(N, dtype=torch.float64).pin_memory()
stream = torch.cuda.Stream()


def bw(reps=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record()
        for _ in range(reps):
            host.copy_(dev, non_blocking=True)
        e1.record()
    e1.synchronize()
    return N * 8 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


bw(3)
alone = 0.0
for r in range(world):
    barrier()
    if r == rank:
        alone = bw()
barrier()
together = bw()
barrier()
vals = torch.tensor([alone, together], dtype=torch.float64, device="cuda")
if world > 1:
    allv = [torch.zeros_like(vals) for _ in range(world)]
    dist.all_gather(allv, vals)
else:
    allv = [vals]
if rank == 0:
    a = [float(v[0]) for v in allv]
    t = [float(v[1]) for v in allv]
    print(json.dumps({"ranks": world, "bytes": N * 8, "alone_GBps": [round(x, 1) for x in a], "together_GBps": [round(x, 1) for x in t],
                      "aggregate_together_GBps": round(sum(t), 1), "host_cores": os.cpu_count()}))
if world > 1:
    dist.destroy_process_group()
