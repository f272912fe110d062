"""Aggregate host<->device copy bandwidth with all ranks copying at once (run under torchrun): the ceiling of the e2e
number at N GPUs.  Each rank copies a 256 MB pinned buffer to its GPU (and back) 20 times between barriers; prints one
JSON line on rank 0 with the per-rank and aggregate GB/s, with every rank copying at once (run it with one rank for the single-GPU figure)."""
import json, os, time
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
N = 256 << 20
h = torch.empty(N, dtype=torch.uint8).pin_memory(); h.fill_(1)
d = torch.empty(N, dtype=torch.uint8, device=dev)
h2 = torch.empty(N, dtype=torch.uint8).pin_memory()
d2 = torch.empty(N, dtype=torch.uint8, device=dev)
s2 = torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def run(direction, reps=20):
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if direction == "h2d":
            d.copy_(h, non_blocking=True)
        elif direction == "d2h":
            h2.copy_(d, non_blocking=True)
        else:                                    # both directions at once, on two streams (GB/s is per direction)
            d.copy_(h, non_blocking=True)
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return N * reps / dt / 1e9


out = {}
for direction in ("h2d", "d2h", "both"):
    run(direction, 3)
    together = run(direction)
    t = torch.tensor([together], dtype=torch.float64, device=dev)
    g = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(g, t)
    else:
        g = [t]
    out[direction] = {"per_rank_GBps_all_ranks_copying": [round(float(x), 1) for x in g], "aggregate_GBps": round(float(sum(float(x) for x in g)), 1)}
if rank == 0:
    print(json.dumps({"n_gpus": world, "buffer_MB": N >> 20, **out}), flush=True)
if world > 1:
    dist.destroy_process_group()
