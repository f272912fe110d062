"""N>1 host logic on CPU (gloo, world_size 2): the read partition covers every read exactly once, per-rank
results merge back into input order, and a sharded run of the path equals the unsharded one.  The per-rank
compute stand-in here is the oracle (test infrastructure): the GPU arm of the same flow is bench.py under torchrun."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_reads_partition_properties():
    from rawalign_b200.shard import shard_reads
    rng = np.random.default_rng(0)
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 5, 100, 1000):
            lens = rng.integers(1, 120_000, n)
            parts = shard_reads(lens, world)
            assert len(parts) == world
            assert parts[0][0] == 0 and parts[-1][1] == n
            for a, b in zip(parts[:-1], parts[1:]):
                assert a[1] == b[0] and a[0] <= a[1]
            if n >= 100 * world:          # balanced within one read of the ideal split
                loads = [int(lens[lo:hi].sum()) for lo, hi in parts]
                assert max(loads) - min(loads) <= 2 * int(lens.max())


def _worker(rank, world, port, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    os.environ["RANK"] = str(rank); os.environ["WORLD_SIZE"] = str(world)
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import Oracle
    from workload import synth
    from rawalign_b200.shard import merge_in_order, shard_reads
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ref = synth.random_reference(20_000, seed=1)
        reads = synth.simulate_reads(ref, 12, seed=9, max_chunks=2)      # same seed on every rank: same input batch
        lo, hi = shard_reads(reads.read_len, world)[rank]
        orc = Oracle()
        rows = []
        for r in range(lo, hi):
            coff, clen, _ = synth.chunk_table(reads.read_off[r:r + 2], reads.read_len[r:r + 1], max_chunks=2)
            ev, off = orc.detect_events_batch(reads.sig, coff, clen)
            rows.append((r, ev.tobytes(), off.tolist()))
        merged = merge_in_order(rows, (lo, hi), len(reads.read_len))
        if rank == 0:
            assert [m[0] for m in merged] == list(range(len(reads.read_len)))
            for r, evb, off in merged:                                     # equals the unsharded computation
                coff, clen, _ = synth.chunk_table(reads.read_off[r:r + 2], reads.read_len[r:r + 1], max_chunks=2)
                ev, off2 = orc.detect_events_batch(reads.sig, coff, clen)
                assert ev.tobytes() == evb and off2.tolist() == off
            open(os.path.join(tmp, "ok"), "w").write("ok")
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_equals_unsharded_world2(tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()


def test_broadcast_index_world1_is_a_plain_load_signature():
    """broadcast_index / Index.alloc_like / Index.tables exist with the documented signatures (the collective itself needs GPUs:
    it is exercised by bench.py under torchrun and by tests/test_gpu_map.py::test_index_replica_by_copy on one GPU)."""
    import inspect
    from rawalign_b200 import shard
    from rawalign_b200.mapper import Index
    assert list(inspect.signature(shard.broadcast_index).parameters) == ["ctx", "index_path", "rank", "device"]
    assert hasattr(Index, "alloc_like") and hasattr(Index, "tables") and hasattr(Index, "counts")
