"""Debug helper: chain-score parity of the bench workload at a given batch size (GPU box)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import rawalign_b200 as ra
from oracle import DtwOpt, Oracle

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ref_seq, ref_arena = bench.build_reference()
reads, coff, clen, crid = bench.build_reads(ref_seq, n_reads, seed=2)
ctx = ra.Context(0); ctx.ref_load_flat(ref_arena)
t0 = time.perf_counter()
ev, off = ctx.detect_events_batch(reads.sig, coff, clen)
t1 = time.perf_counter()
ev_pos = ctx.last_event_positions(len(ev))
anchors, chains = bench.build_chains(off, ev_pos, coff, crid, reads, n_reads, bench.REF_BASES - 6, seed=3)
t2 = time.perf_counter()
score, keep, issued, useful = ctx.align_chains_batch(None, None, anchors, chains, min_score_mode=1, count_cells=True)
t3 = time.perf_counter()
print(f"events {1e3*(t1-t0):.1f} ms  chains {1e3*(t3-t2):.1f} ms  n_chains={len(chains)} issued={issued} useful={useful}")
orc = Oracle(); oo = DtwOpt.default()
rng = np.random.default_rng(0)
groups = np.unique(chains["group"])
pick = np.concatenate([groups[:16], rng.choice(groups, 48, replace=False), groups[-16:]])
bad = 0
for g in pick:
    sel = np.nonzero(chains["group"] == g)[0]
    best = 0.0
    for k in sel:
        ch = chains[k]
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        s = orc.align_chain(a, ev[int(ch["read_off"]):], ref_arena[int(ch["ref_off"]):], oo, min_score=best, score_fma=1)
        if np.float32(s).view(np.uint32) != score[k].view(np.uint32):
            bad += 1
            if bad < 10: print("MISMATCH group", g, "chain", k, "gpu", score[k], "oracle", s, "n_anchors", ch["n_anchors"], "best", best)
        if s >= 20.0 and s > best: best = float(s)
print("mismatches:", bad, "of groups", len(pick))
# timing of e2e pieces
import torch
raw_h = torch.from_numpy(reads.raw).pin_memory().numpy(); ev_host = torch.empty(len(ev) + 1024, dtype=torch.float32).pin_memory().numpy()
anchors_p = torch.from_numpy(anchors.view(np.uint8).reshape(-1)).pin_memory().numpy().view(anchors.dtype)
chains_p = torch.from_numpy(chains.view(np.uint8).reshape(-1)).pin_memory().numpy().view(chains.dtype)
for it in range(4):
    t0 = time.perf_counter()
    e2, o2, ls, fi = ctx.detect_events_raw_batch(raw_h, reads.read_off[:-1], reads.read_len, reads.dac_offset, reads.dac_scale, 4000, bench.MAX_CHUNKS, ev_host)
    t1 = time.perf_counter()
    s2, k2, _, _ = ctx.align_chains_batch(None, None, anchors_p, chains_p, min_score_mode=1, count_cells=False)
    t2 = time.perf_counter()
    print(f"e2e: raw events {1e3*(t1-t0):.2f} ms  chains {1e3*(t2-t1):.2f} ms  equal={np.array_equal(e2, ev)} {np.array_equal(s2.view(np.uint32), score.view(np.uint32))}")
# raw PCIe numbers of this box (pinned)
dev = torch.device("cuda", 0)
src = torch.empty(260_000_000, dtype=torch.uint8).pin_memory(); dst = torch.empty_like(src, device=dev)
back = torch.empty(60_000_000, dtype=torch.uint8).pin_memory(); dsrc = torch.empty(60_000_000, dtype=torch.uint8, device=dev)
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); dst.copy_(src, non_blocking=True); torch.cuda.synchronize(); t1 = time.perf_counter()
    back.copy_(dsrc, non_blocking=True); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"H2D 260 MB {1e3*(t1-t0):.2f} ms = {0.26/(t1-t0):.1f} GB/s   D2H 60 MB {1e3*(t2-t1):.2f} ms = {0.06/(t2-t1):.1f} GB/s")
