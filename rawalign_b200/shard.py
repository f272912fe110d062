"""Multi-GPU partitioning of the hot path: reads are independent, so an input batch is cut
into ``world`` contiguous read ranges balanced by total signal length; rank g runs event detection and every
DTW round for its range.  The only replicated state is the reference-signal mirror, broadcast once from
rank 0; the hot path itself has NO collective.  Results are merged on the host by read index, i.e. in input
order, which is what the reference's output loop relies on (src/rmap.cpp:954-972).

One process per GPU (torchrun); ``torch.distributed`` is plumbing only (nccl on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def shard_reads(read_len: Sequence[int], world: int) -> List[Tuple[int, int]]:
    """Contiguous [lo, hi) read ranges, one per rank, balanced by the number of samples the mapper will look at.
    Every read belongs to exactly one range; ranges are in rank order and cover [0, n_reads)."""
    w = np.asarray(read_len, np.int64)
    n = len(w)
    if world <= 0:
        raise ValueError("world must be positive")
    csum = np.concatenate(([0], np.cumsum(w)))
    total = int(csum[-1])
    cuts = [0]
    for g in range(1, world):
        # first read index whose prefix reaches g/world of the samples; never before the previous cut
        target = total * g / world
        k = int(np.searchsorted(csum, target, "left"))
        cuts.append(min(n, max(k, cuts[-1])))
    cuts.append(n)
    return [(cuts[g], cuts[g + 1]) for g in range(world)]


def broadcast_reference(ctx, ref_arena, rank: int, device=None) -> None:
    """Replicates the reference-signal mirror: rank 0 holds ``ref_arena`` (host float32), every rank ends up with
    it in the context's device arena.  One broadcast over NCCL (NVLink/NVSwitch); world size 1 is a plain load."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        ctx.ref_load_flat(ref_arena)
        return
    n = torch.tensor([len(ref_arena) if rank == 0 else 0], dtype=torch.int64, device=device)
    dist.broadcast(n, 0)
    n = int(n.item())
    ctx.ref_alloc(n)
    ptr, _ = ctx.ref_device()

    class _Mirror:  # the context's device arena viewed as a tensor (no copy)
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}
    mirror = torch.as_tensor(_Mirror(), device=device)
    if rank == 0:
        mirror.copy_(torch.from_numpy(np.ascontiguousarray(ref_arena, np.float32)))
    dist.broadcast(mirror, 0)
    torch.cuda.synchronize(device)


def broadcast_index(ctx, index_path, rank: int, device=None):
    """Replicates the index: rank 0 parses the reference's index file (ra_index_load_file), every other rank allocates an index
    of the same geometry and receives the three device tables -- hash slots, positions, reference signals -- by one broadcast
    each over NCCL (NVLink / NVSwitch).  Returns the rank's Index.  World size 1 is a plain load."""
    import torch
    import torch.distributed as dist
    from .mapper import Index
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return Index.load_file(ctx, index_path)
    meta = [None]
    ix = None
    if rank == 0:
        ix = Index.load_file(ctx, index_path)
        meta = [dict(info=ix.counts(), seq_len=ix.seq_len.tolist(), names=ix.names, bytes=[b for _, b in ix.tables()])]
    dist.broadcast_object_list(meta, 0)
    m = meta[0]
    if rank != 0:
        # the replica needs the table geometry, not the exact key count: 16-byte slots, 2^bits >= 2 * n_keys + 16 of them
        n_vals = m["bytes"][1] // 8
        n_keys = (m["bytes"][0] // 16 - 16) // 2
        ix = Index.alloc_like(ctx, m["info"], n_keys, n_vals, np.array(m["seq_len"], np.uint32), m["names"])
        assert [b for _, b in ix.tables()] == m["bytes"], "replica geometry differs from rank 0's"
    for ptr, nbytes in ix.tables():
        if nbytes == 0:
            continue

        class _T:
            __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        t = torch.as_tensor(_T(), device=device)
        dist.broadcast(t, 0)
    torch.cuda.synchronize(device)
    return ix


def merge_in_order(local_rows: list, my_range: Tuple[int, int], n_reads: int, dst: int = 0):
    """Gathers per-read results (any picklable rows, one per read of ``my_range``) on rank ``dst`` and returns the
    list of all ``n_reads`` rows in input order there (None elsewhere).  Host-side, after the GPU work."""
    import torch.distributed as dist
    lo, hi = my_range
    if len(local_rows) != hi - lo:
        raise ValueError(f"rank holds {len(local_rows)} rows for reads [{lo},{hi})")
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return list(local_rows)
    rank, world = dist.get_rank(), dist.get_world_size()
    gathered = [None] * world if rank == dst else None
    dist.gather_object((lo, hi, local_rows), gathered, dst=dst)
    if rank != dst:
        return None
    out = [None] * n_reads
    for lo_g, hi_g, rows in gathered:
        out[lo_g:hi_g] = rows
    if any(r is None for r in out):
        raise RuntimeError("merge_in_order: some reads were not covered by any rank")
    return out
