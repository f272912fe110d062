#!/usr/bin/env python
"""bench.py -- RawAlign hot path (event detection + DTW chain evaluation) on B200.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA, through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code for the path
                                                           (oracle/_ref when built, else the oracle port)

Workload (BASELINE.json configs[1], synthetic): 5 Mb random reference, pore-model-simulated R9.4 reads,
sparse border, banded=0.1.  One step = one pass of the hot path over one batch of reads per GPU:
K1 (event detection) over every 4000-sample chunk of the batch, then the DTW evaluation of every
candidate chain of every chunk round (chains grow with the number of chunks seen and are recomputed
from scratch each round, as rmap.cpp does).  Weak scaling: every rank gets its own batch (reads are
independent; no data-path collective).  See DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# two batches are in flight per GPU (PIPELINE_DEPTH contexts, 8 streams each): give every stream its own hardware queue
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
PIPELINE_DEPTH = int(os.environ.get("RA_BENCH_PIPELINE", 2))
# measured with ncu on this build (profiles/r01_ncu_full_v5_k1.ncu-rep, profiles/r01_k1_v5_summary.md): DRAM bytes and warp
# instructions of the K1 kernels per signal sample
K1_DRAM_BYTES_PER_SAMPLE = 35.75
K1_WARP_INSTR_PER_SAMPLE = 10.64

REF_BASES = 5_000_000
MAX_CHUNKS = 8
CHAINS_PER_ROUND = 3


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------------
# workload
# ------------------------------------------------------------------------------------------------------
def build_reference():
    from workload import synth
    ref = synth.random_reference(REF_BASES, seed=1)
    mm, _ = synth.pore_model()
    fwd = synth.seq_to_sig(ref, mm, 0); rev = synth.seq_to_sig(ref, mm, 1)
    return ref, np.concatenate([fwd, rev])          # arena: [strand0 | strand1]


def build_reads(ref, n_reads, seed):
    from workload import synth
    reads = synth.simulate_reads(ref, n_reads, seed=seed, max_chunks=MAX_CHUNKS)
    coff, clen, crid = synth.chunk_table(reads.read_off, reads.read_len, max_chunks=MAX_CHUNKS)
    return reads, coff, clen, crid


def build_chains(ev_off, ev_pos, coff, crid, reads, n_reads, ref_len, seed):
    """Chains of every chunk round of every read.  Round c of a read sees the events of chunks 0..c (contiguous in
    the CSR event arena because chunks of a read are consecutive).  Per round: the true chain (anchors placed from the
    simulator's ground truth, as exact seed hits would be) followed by CHAINS_PER_ROUND-1 decoys at random positions,
    i.e. the reference's evaluation order, best chaining score first."""
    from rawalign_b200 import CHAIN_DTYPE
    from workload import synth
    rng = np.random.default_rng(seed)
    per_chunk = np.diff(ev_off.astype(np.int64))
    first_chunk = np.searchsorted(crid, np.arange(n_reads), "left")
    last_chunk = np.searchsorted(crid, np.arange(n_reads), "right")
    half = ref_len + 6 - 5                                  # strand-1 signal follows strand-0 in the arena (len = bases-5 each)
    anchors, chains = [], []
    a_base = 0
    for rnd in range(MAX_CHUNKS):
        for r in range(n_reads):
            nch = last_chunk[r] - first_chunk[r]
            if nch <= rnd:
                continue
            c0 = first_chunk[r]
            e0 = int(ev_off[c0]); e1 = int(ev_off[c0 + rnd + 1])
            ne = e1 - e0
            if ne < 60:
                continue
            # sample position of every event within the read
            chunk_of = np.repeat(np.arange(rnd + 1), per_chunk[c0:c0 + rnd + 1])
            spos = ev_pos[e0:e1].astype(np.int64) + chunk_of * synth.CHUNK_SIZE
            ref_off = 0 if reads.strand[r] == 0 else half
            grp = []
            a = synth.truth_chain(spos, reads.base_start[r], int(reads.ref_start[r]), rng)
            if a is not None:
                grp.append((a, ref_off))
            da, dc = synth.synth_chains(np.array([ne]), np.array([e0], np.uint64), ref_len - 8, seed=int(rng.integers(1 << 30)),
                                        chains_per_read=CHAINS_PER_ROUND)
            for ch in dc[1:]:
                grp.append((da[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])], int(rng.integers(0, 2)) * half))
            for a, roff in grp:
                ch = np.zeros(1, CHAIN_DTYPE)
                ch["anchor_off"] = a_base; ch["n_anchors"] = len(a); ch["group"] = r * MAX_CHUNKS + rnd
                ch["read_off"] = e0; ch["ref_off"] = roff
                anchors.append(a); chains.append(ch); a_base += len(a)
    return np.concatenate(anchors), np.concatenate(chains)


def chains_to_items(anchors, chains, frac=0.1, return_chain_of=False):
    """Sub-problems of sparse-border align_chain (rmap.cpp:248-281), vectorised; test/bench helper."""
    from rawalign_b200 import ITEM_DTYPE
    n_parts = chains["n_anchors"].astype(np.int64) - 1
    tot = int(n_parts.sum())
    chain_of = np.repeat(np.arange(len(chains)), n_parts)
    first = np.concatenate(([0], np.cumsum(n_parts)[:-1]))
    p = np.arange(tot) - first[chain_of]
    parts = n_parts[chain_of]
    s_idx = chains["anchor_off"].astype(np.int64)[chain_of] + parts - p
    s = anchors[s_idx]; e = anchors[s_idx - 1]
    it = np.zeros(tot, ITEM_DTYPE)
    it["a_off"] = chains["read_off"][chain_of] + s["query_position"]
    it["b_off"] = chains["ref_off"][chain_of] + s["target_position"]
    it["a_len"] = e["query_position"] - s["query_position"] + 1
    it["b_len"] = e["target_position"] - s["target_position"] + 1
    it["band_radius"] = np.maximum(1, (it["a_len"].astype(np.float32) * np.float32(frac)).astype(np.int32))
    it["exclude_last"] = (p != parts - 1)
    if return_chain_of:
        return it, chain_of
    return it


def item_cells(items):
    """Exact DP cells per item (GCUPS numerator) -- closed form for the shapes that dominate, C ABI otherwise."""
    import rawalign_b200 as ra
    key = (items["a_len"].astype(np.uint64) << np.uint64(40)) | (items["b_len"].astype(np.uint64) << np.uint64(16)) | items["band_radius"].astype(np.uint64)
    uniq, inv = np.unique(key, return_inverse=True)
    cells = np.array([ra.Context.dtw_cells(int(k >> np.uint64(40)), int((k >> np.uint64(16)) & np.uint64(0xFFFFFF)), int(k & np.uint64(0xFFFF))) for k in uniq], np.int64)
    return cells[inv]


# ------------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed regions: NVML in-process (about 1 ms per sample), nvidia-smi as
    the fallback."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index; self.sm = []; self.mx = []; self.reasons = set(); self._stop = False; self.t = None; self.how = "nvml"

    def _nvml_loop(self):
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
        h = nv.nvmlDeviceGetHandleByIndex(idx)
        self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)))
        bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop:
            self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            for name, bit in bits.items():
                if r & bit:
                    self.reasons.add(name)
            time.sleep(0.002)

    def _smi_loop(self):
        self.how = "nvidia-smi"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                r = [x.strip() for x in out.split(",")]
                if r and r[0].replace(".", "").isdigit():
                    self.sm.append(float(r[0])); self.mx.append(float(r[1]))
                    for i in range(4):
                        if len(r) > 2 + i and r[2 + i].lower().startswith("active"):
                            self.reasons.add(names[i])
            except Exception:
                pass
            time.sleep(0.05)

    def start(self):
        def run():
            try:
                self._nvml_loop()
            except Exception:
                self._smi_loop()
        self.t = threading.Thread(target=run, daemon=True); self.t.start()

    def stop(self):
        self._stop = True
        if self.t:
            self.t.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "how": self.how}


# ------------------------------------------------------------------------------------------------------
# CPU legs
# ------------------------------------------------------------------------------------------------------
def cpu_backend():
    """The reference's own compiled map-loop code (oracle/_ref/libra_refmap.so: unmodified rmap.cpp/revent.c/dtw.cpp behind
    refmap_shim.cpp) when it was built, else the C restatement (oracle port)."""
    from oracle import RefMap
    if RefMap.available():
        return RefMap(), "reference"
    from oracle import Oracle
    return Oracle(), "port"


def read_major(chains, n_reads):
    """bench chains are round-major (one ra_align_chains_batch call covers a round of every read); the reference's map loop
    is read-major (rmap.cpp:685-693).  Returns the chains re-ordered per read and the (n_reads, MAX_CHUNKS+1) table of chain
    ranges per chunk round that refmap_hotpath_batch consumes."""
    order = np.argsort(chains["group"], kind="stable")
    ch = chains[order]
    g = ch["group"].astype(np.int64)
    counts = np.bincount(g, minlength=n_reads * MAX_CHUNKS).reshape(n_reads, MAX_CHUNKS)
    rf = np.zeros((n_reads, MAX_CHUNKS + 1), np.int64)
    rf[:, 1:] = np.cumsum(counts, axis=1)
    rf += np.concatenate(([0], np.cumsum(counts.sum(axis=1))[:-1]))[:, None]
    return ch, rf.astype(np.uint32), order


def cpu_step(be, kind, reads, n_reads, anchors, chains_rm, round_first, ref_arena, n_threads, events_only=False):
    """One pass of the hot path on the host: ONE C call.  kind 'reference': refmap_hotpath_batch runs, per read on a
    dynamically scheduled thread (as kt_for does), detect_events per chunk (rmap.cpp:551), the event append (:554-567) and
    the reference's own align_chain over the chains of that chunk round with the running best_found_alignment
    (:515-524) -- budget pre-check and early termination are the reference's.  Nothing else is inside the timer."""
    t0 = time.perf_counter()
    if kind == "reference":
        opt = be.mapopt()
        tot, score, keep, cnt, _, _ = be.hotpath_batch(reads.sig, reads.read_off[:n_reads], reads.read_len[:n_reads], ref_arena,
                                                       None if events_only else anchors, None if events_only else chains_rm,
                                                       None if events_only else round_first[:n_reads], opt, n_threads=n_threads, max_chunks=MAX_CHUNKS)
    else:
        tot, score = port_step(be, reads, n_reads, anchors, chains_rm, round_first, ref_arena, events_only)
    return time.perf_counter() - t0, tot, score


def port_step(orc, reads, n_reads, anchors, chains_rm, round_first, ref_arena, events_only):
    """Fallback when oracle/_ref is absent: the same loop on the C restatement, single thread."""
    from oracle import DtwOpt
    from workload import synth
    coff, clen, crid = synth.chunk_table(reads.read_off[:n_reads + 1], reads.read_len[:n_reads], max_chunks=MAX_CHUNKS)
    ev, off = orc.detect_events_batch(reads.sig, coff, clen)
    score = np.zeros(len(chains_rm), np.float32)
    if not events_only:
        oo = DtwOpt.default()
        first_chunk = np.searchsorted(crid, np.arange(n_reads), "left")
        for r in range(n_reads):
            e0 = int(off[first_chunk[r]])
            for c in range(MAX_CHUNKS):
                best = 0.0
                for k in range(int(round_first[r, c]), int(round_first[r, c + 1])):
                    ch = chains_rm[k]
                    a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
                    s = orc.align_chain(a, ev[e0:], ref_arena[int(ch["ref_off"]):], oo, min_score=best, score_fma=1)
                    score[k] = s
                    if s >= 20.0 and s > best:
                        best = float(s)
    return len(ev), score


def useful_cells(chains_rm, anchors, score, frac=0.1):
    """DP cells the reference evaluated (the GCUPS numerator): chains dropped by the budget pre-check before
    their first part cost nothing; for the others every part up to the drop.  Computed from the scores the reference arm
    returned plus the C ABI's exact per-item cell count -- outside the timed region."""
    import rawalign_b200 as ra
    items, chain_of = chains_to_items(anchors, chains_rm, frac, return_chain_of=True)
    cells = item_cells(items)
    alive = score[chain_of] != np.float32(-1e10)
    return int(cells[alive].sum()), int(cells.sum())


# ------------------------------------------------------------------------------------------------------
def bind_to_gpu_numa_node(index):
    """Pin this process to the CPUs NVML reports as local to GPU `index` (the cores of the socket / NUMA node its PCIe root
    hangs off), BEFORE CUDA initialises and before any pinned buffer is allocated: pinned pages are placed on the node of
    the allocating thread, and a host<->device copy whose pages sit on the other socket crosses the inter-socket link,
    which eight ranks share.  Returns a short description for the JSON line; never fatal."""
    if os.environ.get("RA_BENCH_NO_BIND"):
        return "off (RA_BENCH_NO_BIND)"
    try:
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
        h = nv.nvmlDeviceGetHandleByIndex(idx)
        n_cpu = os.cpu_count() or 64
        words = nv.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "unchanged (NVML reports no local CPUs within the allowed set)"
        os.sched_setaffinity(0, cpus)
        return f"gpu-local: {len(cpus)} of {n_cpu} cpus [{min(cpus)}..{max(cpus)}]"
    except Exception as e:   # noqa: BLE001 -- a box without NVML affinity data still runs the bench
        return f"unchanged ({type(e).__name__}: {e})"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5", "C2hot"],
                    help="BASELINE.json configs through the whole map loop (ra_map_batch); C2hot = only the dense hot-path step "
                         "(K1 + chain DTW on synthetic chains) that also supplies the roofline numbers of the default line")
    ap.add_argument("--reads-per-gpu", type=int, default=int(os.environ.get("RA_BENCH_READS", 4096)))
    ap.add_argument("--map-reads-per-gpu", type=int, default=int(os.environ.get("RA_BENCH_MAP_READS", 8192)))
    ap.add_argument("--cpu-sample-reads", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-hot-path", action="store_true", help="skip the dense hot-path step in a map-loop run (no roofline numbers)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 3)
    hot_config = {"workload": "C2 hot path: 5 Mb synthetic reference, simulated R9.4 reads, sparse border, banded=0.1; K1 over every chunk + chain DTW of every chunk round on synthetic chains",
                  "reads_per_gpu": args.reads_per_gpu, "batch_note": "one batch ~ the reference's default mini-batch (-K 500M of signal, roptions.c:34)",
                  "max_chunks_per_read": MAX_CHUNKS, "chains_per_round": CHAINS_PER_ROUND, "batches_in_flight_per_gpu": max(1, PIPELINE_DEPTH),
                  "l2_policy": "inputs larger than L2 (signal batch >> 126 MB)"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        if args.config == "C2hot":
            return run_reference_arm(args, hot_config)
        return run_reference_map_arm(args)

    # exactly one line on stdout: anything a C library prints (e.g. the NCCL version banner) goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1); os.dup2(2, 1)
    binding = bind_to_gpu_numa_node(local)
    hot_config["cpu_binding"] = binding
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback in the product path); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    if args.config == "C2hot":
        out = hot_path(args, hot_config, rank, world, local, torch, dist, dev)
    else:
        out = map_path(args, rank, world, local, torch, dist, dev, binding)
        if not args.no_hot_path:
            hot = hot_path(args, hot_config, rank, world, local, torch, dist, dev, with_cpu=False)
            if rank == 0:
                out["roofline"] = hot["roofline"]; out["roofline_dtw"] = hot["roofline_dtw"]
                out["hot_path"] = {k: hot[k] for k in ("value", "unit", "ms_per_step", "dtw_gcups", "hot_path_gcups", "cells_per_step", "stages_ms", "e2e", "gpu_launches", "config")}
                out["hot_path"]["what"] = ("the dense form of the same path (--config C2hot): K1 over every chunk of 4096 reads x 8 chunks and the chain DTW of every "
                                           "chunk round on synthetic chains, no early stop -- the step the roofline numbers are measured on")
                out["gpu_launches"] += hot["gpu_launches"]
    if rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


def hot_path(args, config, rank, world, local, torch, dist, dev, with_cpu=True):
    """The dense hot-path step (K1 + chain DTW, synthetic chains); returns the JSON dict on rank 0."""
    import rawalign_b200 as ra
    out = None
    t_setup = time.time()
    ref_seq, ref_arena = build_reference()
    reads, coff, clen, crid = build_reads(ref_seq, args.reads_per_gpu, seed=2 + rank)
    hook_opt = ra.ri_mapopt_init(); hook_opt.reserved = int(os.environ.get("RA_BENCH_RESERVED", "0"))   # A/B of kernel variants (test hooks)
    ctx = ra.Context(local, hook_opt)
    # reference mirror: rank 0 loads, everyone else receives it by one NCCL broadcast
    from rawalign_b200.shard import broadcast_reference
    broadcast_reference(ctx, ref_arena if rank == 0 else None, rank, dev)
    # software pipelining across steps: PIPELINE_DEPTH contexts on this GPU, each driven by its own host thread, so that one
    # batch's uploads / latency-bound kernels overlap the other's issue-bound kernels (what a production map loop does with
    # double-buffered batches).  Every step still does all of its work.
    ctxs = [ctx]
    for _ in range(1, max(1, PIPELINE_DEPTH)):
        c2 = ra.Context(local, hook_opt)
        c2.ref_alloc(len(ref_arena))
        p_src, _n = ctx.ref_device(); p_dst, _n = c2.ref_device()
        _dev_copy(torch, dev, p_dst, p_src, len(ref_arena))
        ctxs.append(c2)

    n_chunks = len(coff)
    sig_pinned = torch.from_numpy(reads.sig).pin_memory()
    coff_p = torch.from_numpy(coff.view(np.int64)).pin_memory(); clen_p = torch.from_numpy(clen.view(np.int32)).pin_memory()
    sig_d = sig_pinned.to(dev); coff_d = coff_p.to(dev); clen_d = clen_p.to(dev)
    ev_cap = int(clen.sum()) // 2 + n_chunks + 16
    ev_d = torch.empty(ev_cap, dtype=torch.float32, device=dev)
    evoff_d = torch.empty(n_chunks + 1, dtype=torch.int64, device=dev)
    max_len = int(clen.max())
    torch.cuda.synchronize()

    ev_ds = [ev_d] + [torch.empty_like(ev_d) for _ in ctxs[1:]]
    evoff_ds = [evoff_d] + [torch.empty_like(evoff_d) for _ in ctxs[1:]]

    def k1_dev(w=0):
        return ctxs[w].detect_events_batch_dev(sig_d.data_ptr(), coff_d.data_ptr(), clen_d.data_ptr(), n_chunks, max_len,
                                               ev_ds[w].data_ptr(), ev_ds[w].numel(), evoff_ds[w].data_ptr())
    n_events = k1_dev()
    ev_off = evoff_d.cpu().numpy().view(np.uint64)
    ev_pos = ctx.last_event_positions(n_events)
    anchors, chains = build_chains(ev_off, ev_pos, coff, crid, reads, args.reads_per_gpu, REF_BASES - 6, seed=3 + rank)
    gfirst = np.concatenate((np.nonzero(np.concatenate(([True], chains["group"][1:] != chains["group"][:-1])))[0], [len(chains)])).astype(np.uint32)
    n_groups = len(gfirst) - 1
    # pinned host copies (the e2e path copies from these every step) and device-resident copies (the `value` path)
    anchors_p = torch.from_numpy(anchors.view(np.uint8).reshape(-1)).pin_memory()
    chains_p = torch.from_numpy(chains.view(np.uint8).reshape(-1)).pin_memory()
    anchors_d = anchors_p.to(dev); chains_d = chains_p.to(dev); gfirst_d = torch.from_numpy(gfirst.view(np.int32)).to(dev)
    score_d = torch.empty(len(chains), dtype=torch.float32, device=dev); keep_d = torch.empty(len(chains), dtype=torch.uint8, device=dev)
    anchors_h = anchors_p.numpy().view(anchors.dtype); chains_h = chains_p.numpy().view(chains.dtype)
    n_parts = int((chains["n_anchors"].astype(np.int64) - 1).sum())
    n_samples = int(clen.sum())

    score_ds = [score_d] + [torch.empty_like(score_d) for _ in ctxs[1:]]
    keep_ds = [keep_d] + [torch.empty_like(keep_d) for _ in ctxs[1:]]

    def k2_dev(count=False, w=0):
        return ctxs[w].align_chains_batch_dev(ev_ds[w].data_ptr(), None, anchors_d.data_ptr(), len(anchors), chains_d.data_ptr(), len(chains), gfirst_d.data_ptr(),
                                              n_groups, score_ds[w].data_ptr(), keep_ds[w].data_ptr(), min_score_mode=1, count_cells=count)
    cells_issued, cells_useful = k2_dev(True)      # cell counting is instrumentation: one counted pass, timed passes run without it
    log(f"[rank {rank}] setup {time.time() - t_setup:.1f}s: reads={args.reads_per_gpu} chunks={n_chunks} samples={n_samples} "
        f"events={n_events} chains={len(chains)} groups={n_groups} sub-DTWs={n_parts} cells issued={cells_issued} useful={cells_useful}")

    ev_hosts = [torch.empty(ev_cap, dtype=torch.float32).pin_memory().numpy() for _ in ctxs]
    raw_h = torch.from_numpy(reads.raw).pin_memory().numpy()
    raw_off = np.ascontiguousarray(reads.read_off[:-1]); raw_len = reads.read_len
    e2e_out = {}

    def step_e2e(w=0):
        # int16 DAC samples (what a FAST5 holds) -> device in slices that overlap the kernels, decode + K1, events back to the
        # host (the map loop seeds from them), anchors/chains -> device, chain evaluation against the events that stayed on the
        # device, scores back
        ev, off, lsig, first = ctxs[w].detect_events_raw_batch(raw_h, raw_off, raw_len, reads.dac_offset, reads.dac_scale, 4000, MAX_CHUNKS, ev_hosts[w])
        s, keep, _, _ = ctxs[w].align_chains_batch(None, None, anchors_h, chains_h, min_score_mode=1, count_cells=False)
        if w == 0:
            e2e_out["score"] = s; e2e_out["n_ev"] = len(ev); e2e_out["ev"] = ev

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def run_steps(step_fn, n_steps):
        """n_steps steps dealt round-robin to the pipelines (one host thread per context); returns when all are done."""
        if len(ctxs) == 1:
            for _ in range(n_steps):
                step_fn(0)
            return
        errs = []

        def worker(w):
            try:
                for _ in range(w, n_steps, len(ctxs)):
                    step_fn(w)
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        ths = [threading.Thread(target=worker, args=(w,)) for w in range(len(ctxs))]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if errs:
            raise errs[0]

    def step_dev(w=0):
        k1_dev(w); k2_dev(False, w)

    # ---- device-resident timing ----
    run_steps(step_dev, max(args.warmup, len(ctxs)) * len(ctxs))
    barrier()
    # per-stage device times: one un-pipelined step on context 0 (CUDA events on the library streams)
    ev_ms, dtw_ms, det_ms = [], [], []
    for _ in range(3):
        step_dev(0)
        tm = ctx.last_timing()
        det_ms.append(tm[0]); ev_ms.append(tm[1]); dtw_ms.append(tm[2])
    barrier()
    sampler = ClockSampler(local); sampler.start()
    l0 = sum(c.launch_count for c in ctxs)
    t0 = time.perf_counter()
    run_steps(step_dev, args.steps)
    barrier()
    dt = time.perf_counter() - t0
    launches = sum(c.launch_count for c in ctxs) - l0
    # ---- end-to-end timing (host buffers, copies inside) ----
    run_steps(step_e2e, 2 * len(ctxs))
    barrier()
    t0 = time.perf_counter()
    run_steps(step_e2e, args.steps)
    barrier()
    dt_e2e = time.perf_counter() - t0
    clocks = sampler.stop()            # sampled over both timed regions (device-resident and end-to-end)

    tt = torch.tensor([dt, dt_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt, dt_e2e = float(tt[0]), float(tt[1])
    tot = torch.tensor([args.reads_per_gpu, cells_issued, n_samples, n_events], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    reads_all, cells_all, samples_all, events_all = [float(x) for x in tot]

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0)); peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback"
        ms_step = dt / args.steps * 1e3
        k1_ms = float(np.median(ev_ms)); k2_ms = float(np.median(dtw_ms))
        k1_bytes = 4.0 * n_samples + 4.0 * n_events + 8.0 * n_chunks + 12.0 * n_chunks
        alu_peak_gcups = 148 * 128 * 1965.0e6 / 4 / 1e9
        out = {
            "metric": "reads_per_s", "value": reads_all * args.steps / dt, "unit": "reads/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "dtw_gcups": cells_issued / (k2_ms * 1e-3) / 1e9, "hot_path_gcups": cells_all * args.steps / dt / 1e9,
            "cells_per_step": {"issued": cells_issued, "useful": cells_useful},
            "stages_ms": {"events_all_kernels": k1_ms, "events_detect_kernel": float(np.median(det_ms)), "chain_dtw_kernels": k2_ms,
                          "timing": "CUDA events on the library stream, median over the timed steps"},
            "e2e": {"value": reads_all * args.steps / dt_e2e, "unit": "reads/s", "ms_per_step": dt_e2e / args.steps * 1e3,
                    "h2d_bytes_per_step": int(reads.raw.nbytes + 20 * args.reads_per_gpu + coff.nbytes + clen.nbytes + anchors.nbytes + chains.nbytes + gfirst.nbytes),
                    "d2h_bytes_per_step": int(4 * n_events + 8 * (n_chunks + 1) + 4 * args.reads_per_gpu + 5 * len(chains)),
                    "input": "int16 DAC samples, decoded on the device (ra_detect_events_raw_batch); copies sliced to overlap the kernels",
                    "events_equal_resident_path": bool(e2e_out["n_ev"] == n_events and np.array_equal(e2e_out["ev"][:n_events].view(np.uint32), ev_d[:n_events].cpu().numpy().view(np.uint32))),
                    "scores_equal_resident_path": bool(np.array_equal(np.asarray(e2e_out["score"]).view(np.uint32), score_d.cpu().numpy().view(np.uint32)))},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"kernel": "K1 event detection (k_prefix_pipe + k_tstat_pipe + k_fsm_pipe + k_exclusive_scan_u32 + k_finish_pipe, eight chunk groups on concurrent streams)",
                         "bound": "hbm", "achieved": k1_bytes / (k1_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": k1_bytes / (k1_ms * 1e-3) / 1e9 / hbm_peak,
                         "traffic": int(K1_DRAM_BYTES_PER_SAMPLE * n_samples), "peak_source": peak_src,
                         "algorithmic_bytes": "4 B/sample in + 4 B/event out + 20 B/chunk of tables, per step like achieved / traffic",
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of the four K1 kernels, ncu --set full of this build "
                                           "(profiles/r01_ncu_full_v5_k1.ncu-rep): 35.75 B/sample = 4 B signal + the fp32 prefix sums and the two "
                                           "t-stat tracks written once and read once (32 B, part of it absorbed by the 126 MB L2) + the prefix sums re-read at the peaks; the stages are "
                                           "separate kernels, so the intermediates cross HBM -- no stage re-reads its input",
                         "note": "K1 is issue-bound, not HBM-bound: ~10.6 warp instructions per sample (ncu) at "
                                 f"{K1_WARP_INSTR_PER_SAMPLE * n_samples / (k1_ms * 1e-3) / (148 * 4 * clocks.get('sm_mhz', 1965.0) * 1e6):.2f} of the SM issue rate"},
            "roofline_dtw": {"kernel": "sparse chain DTW (k_sp_prep + k_sp_parts2 + k_sp_rest + k_sp_replay, two phases)", "bound": "alu",
                             "achieved": cells_issued / (k2_ms * 1e-3) / 1e9, "peak": alu_peak_gcups,
                             "unit": "GCUPS", "frac": cells_issued / (k2_ms * 1e-3) / 1e9 / alu_peak_gcups,
                             "peak_source": "148 SMs x 128 fp32 lanes x 1.965 GHz / 4 lane-ops per cell (SURVEY 8d)"},
        }
        if with_cpu and not args.no_cpu_baseline:
            be, kind = cpu_backend()
            n_thr = os.cpu_count() or 1
            nsub = min(args.reads_per_gpu, max(args.cpu_sample_reads, 8 * n_thr))
            sel = (chains["group"] // MAX_CHUNKS) < nsub
            ch_rm, rfirst, order = read_major(chains[sel], nsub)
            cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_thr)                         # warm-up (page-in, thread pool)
            t_cpu, n_ev_cpu, sc_cpu = cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_thr)
            t_ev, _, _ = cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_thr, events_only=True)
            t_1, _, _ = cpu_step(be, kind, reads, max(1, nsub // n_thr), anchors, ch_rm, rfirst, ref_arena, 1)
            # dropped chains of this workload are all dropped before their first part (decoys against the truth chain's score)
            cells_ref, cells_all = useful_cells(ch_rm, anchors, sc_cpu)
            gpu_sc = score_d.cpu().numpy()[np.nonzero(sel)[0][order]]
            out["cpu_baseline"] = {"value": nsub / t_cpu, "unit": "reads/s", "cores": n_thr, "kind": kind,
                                   "sample": f"first {nsub} reads of the batch, one pass inside ONE C call (refmap_hotpath_batch: detect_events + the reference's "
                                             f"align_chain with its budget pre-check / early termination, {n_thr} OpenMP threads, dynamic per-read schedule)",
                                   "seconds": t_cpu, "events_only_seconds": t_ev, "one_thread_reads_per_s": max(1, nsub // n_thr) / t_1,
                                   "thread_scaling": (nsub / t_cpu) / (max(1, nsub // n_thr) / t_1),
                                   "dtw_cells_evaluated": cells_ref, "dtw_cells_all_parts": cells_all,
                                   "dtw_gcups": cells_ref / max(t_cpu - t_ev, 1e-9) / 1e9,
                                   "scores_bit_equal_gpu": bool(np.array_equal(sc_cpu.view(np.uint32), gpu_sc.view(np.uint32)))}
            out["cpu_baseline"]["parity_bit_exact"] = out["cpu_baseline"]["scores_bit_equal_gpu"]
    for c in ctxs:
        c.close()
    return out


# ------------------------------------------------------------------------------------------------------
# the whole map loop (ra_map_batch) on the datasets of BASELINE.json's configs
# ------------------------------------------------------------------------------------------------------
def map_opt_kwargs(name):
    from workload.mapbench import CONFIGS
    return dict(CONFIGS[name]["opt"])


def map_path(args, rank, world, local, torch, dist, dev, binding):
    """One step = one batch of reads through the whole map loop on the device: int16 DAC samples in, one PAF record per read out
    (decode, K1, seeding, index lookup, anchor sort, chaining, chain DTW, primary chains / MAPQ / stop test for every chunk
    round).  `value`: the batch already resident in HBM; `e2e`: ra_map_batch from pinned host buffers, copies inside."""
    import rawalign_b200 as ra
    from workload.mapbench import CONFIGS, Dataset, digest
    name = args.config
    cfg = CONFIGS[name]
    if world > 1:                                                 # rank 0 builds the reference and its index file, the others wait and find it cached
        if rank == 0:
            ds = Dataset(name, log=log)
        dist.barrier()
        if rank != 0:
            ds = Dataset(name, log=log)
    else:
        ds = Dataset(name, log=log)
    n_reads = args.map_reads_per_gpu
    t0 = time.time()
    raw, off, ln, dac_off, dac_scale = ds.reads(n_reads, seed=2 + rank)
    log(f"[rank {rank}] {n_reads} reads simulated in {time.time() - t0:.1f}s: {len(raw) * 2 / 1e6:.0f} MB of int16 signal")
    raw_p = torch.from_numpy(raw).pin_memory().numpy()            # the e2e path copies from pinned host memory every step
    # MAP_DEPTH batches in flight per GPU (one context + index each, one host thread each), as a map loop with double-buffered
    # mini-batches runs: one batch's uploads and host round trips overlap the other's kernels.  Every step does all of its work.
    # (a 250 Mb index is 23 GB and its rounds hold ~0.8 G anchors: two batches in flight fit 180 GB, three do not)
    ref_bases = sum(cfg["ref"])
    depth = max(1, int(os.environ.get("RA_BENCH_MAP_DEPTH", 2 if ref_bases >= 200_000_000 else 3 if ref_bases >= 50_000_000 else 4)))
    opt = ra.ri_mapopt_full(**cfg["opt"])
    want_cigar = bool(opt.flag & 0x4)
    pipes = []
    t0 = time.time()
    from rawalign_b200.shard import broadcast_index
    for w in range(depth):
        c = ra.Context(local)
        if os.environ.get("RA_BENCH_HOST_GATHER"):
            c.set_tuning("map_host_gather", int(os.environ["RA_BENCH_HOST_GATHER"]))
        # rank 0 parses the reference's index file; the other ranks receive the device tables (hash slots, positions, reference
        # signals) by NCCL broadcast -- the one collective of the path, at load
        pipes.append((c, broadcast_index(c, ds.index_path, rank, dev) if w == 0 else pipes[0][1].clone(c)))
        if w == 0:
            t_index = time.time() - t0
    ctx, ix = pipes[0]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    cig = {}

    sig_p = torch.from_numpy(ds.decode(raw)).pin_memory().numpy()   # ri_sig_t::sig as the reference's reader hands it to the map loop (rsig.cpp:213-222)

    def step_raw(w=0):
        r = pipes[w][1].map_reads(opt, raw=raw_p, raw_off=off, raw_len=ln, offset=dac_off, scale=dac_scale)
        if want_cigar:
            cig[w] = pipes[w][1].last_cigar(n_reads)             # the aln:s: triples come back to the host inside the timed region
        return r

    def step_e2e(w=0):
        r = pipes[w][1].map_reads(opt, sig=sig_p, sig_off=off, sig_len=ln)
        if want_cigar:
            cig[w] = pipes[w][1].last_cigar(n_reads)
        return r

    def step_res(w=0, count=False):
        r = pipes[w][1].map_reads(opt, resident_reads=n_reads, count_cells=count)
        if want_cigar:
            cig[w] = pipes[w][1].last_cigar(n_reads)
        return r

    last = {}

    def run_steps(fn, n_steps):
        errs = []

        def worker(w):
            try:
                for _ in range(w, n_steps, depth):
                    last[w] = fn(w)
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        ths = [threading.Thread(target=worker, args=(w,)) for w in range(depth)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        if errs:
            raise errs[0]

    for w in range(depth):
        rec0, st0 = step_raw(w)                                   # every pipeline holds the batch; also the first parity sample
    recc, stc = step_res(0, True)                                 # one counted pass (cell counting is instrumentation)
    run_steps(step_res, max(args.warmup, depth))
    barrier()
    # per-stage times of one un-pipelined step
    rec, st_alone = step_res(0)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    l0 = sum(p[0].launch_count for p in pipes)
    t0 = time.perf_counter()
    run_steps(step_res, args.steps)
    barrier()
    dt = time.perf_counter() - t0
    launches = sum(p[0].launch_count for p in pipes) - l0
    rec = last[0][0]
    run_steps(step_raw, depth); barrier()
    t0 = time.perf_counter()
    run_steps(step_raw, args.steps)
    barrier()
    dt_raw = time.perf_counter() - t0
    rec_r, st_r = last[0]
    run_steps(step_e2e, depth); barrier()                         # (the float form leaves nothing resident: it runs last)
    t0 = time.perf_counter()
    run_steps(step_e2e, args.steps)
    barrier()
    dt_e2e = time.perf_counter() - t0
    rec_e, st_e = last[0]
    stage_acc = {k: v * args.steps for k, v in st_alone.items() if k.endswith("_ms")}
    clocks = sampler.stop()
    tt = torch.tensor([dt, dt_e2e, dt_raw], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt, dt_e2e, dt_raw = float(tt[0]), float(tt[1]), float(tt[2])
    tot = torch.tensor([n_reads, stc["cells_useful"], stc["cells_issued"], int(rec0["mapped"].sum())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    reads_all, cells_all, issued_all, mapped_all = [float(x) for x in tot]
    out = None
    if rank == 0:
        lines = ix.paf_lines(rec0, cigar=cig.get(0) if want_cigar else None)
        out = {"metric": "reads_per_s", "value": reads_all * args.steps / dt, "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": cfg["desc"] + "; the whole map loop per read batch (ra_map_batch: decode, event detection, seeding, index lookup, "
                                      "chaining, chain DTW, primary chains, MAPQ, early stop), one step = one batch",
                          "reads_per_gpu": n_reads, "reference_bases": int(sum(cfg["ref"])), "n_sequences": len(cfg["ref"]), "preset": cfg["preset"],
                          "note": cfg.get("note", ""), "index": f"reference index file format, built by {ds.index_by}, loaded by ra_index_load_file on rank 0" + (" and replicated by NCCL broadcast" if world > 1 else "") + f" in {t_index:.2f}s; the other batches in flight use device-to-device copies (ra_index_clone)",
                          "l2_policy": "inputs larger than L2 (signal batch >> 126 MB)", "cpu_binding": binding},
               "reads_mapped_frac": mapped_all / reads_all,
               "dtw_cells_per_step": {"useful": cells_all, "issued": issued_all},
               "hot_path_gcups": cells_all * args.steps / dt / 1e9,
               "map_stages_ms": dict({k: v / args.steps for k, v in sorted(stage_acc.items())}, note="one step run alone (host-observed, between synchronisations); the timed steps overlap the batches in flight"),
               "batches_in_flight_per_gpu": depth,
               "map_counts_per_step": {k: stc[k] for k in ("rounds", "chunks", "events", "seeds", "anchors", "chains_evaluated", "launches")},
               "e2e": {"value": reads_all * args.steps / dt_e2e, "unit": "reads/s", "ms_per_step": dt_e2e / args.steps * 1e3,
                       "h2d_bytes_per_step": int(st_e["h2d_bytes"]),
                       "d2h_bytes_per_step": int(st_e["d2h_bytes"]),   # records + control words + (cigar) 12 B per path element, counted by the library
                       "input": "ri_sig_t of every read (float pA, l_sig) in pinned host memory -- what the reference's reader hands map_worker_for and what "
                                "the reference arm is timed on -> ra_map_batch -> one 88-byte record per read; per chunk round the GPU fetches only the chunk "
                                "each still-active read is at (bytes counted by the library), as the reference's loop touches only those (rmap.cpp:685-712)",
                       "host_buffer_bytes": int(sig_p.nbytes),
                       "stages_ms_last_step": {k: round(float(v), 3) for k, v in st_e.items() if k.endswith("_ms")},
                       "records_equal_resident_path": bool(rec_e.tobytes() == rec.tobytes() == rec0.tobytes())},
               "e2e_int16": {"value": reads_all * args.steps / dt_raw, "unit": "reads/s", "ms_per_step": dt_raw / args.steps * 1e3,
                             "h2d_bytes_per_step": int(st_r["h2d_bytes"]), "d2h_bytes_per_step": int(st_r["d2h_bytes"]),
                             "input": "int16 DAC samples of the whole reads in pinned host memory (decode and the (30,200) pA filter on the device, row f3): "
                                      "l_sig is only known after the whole read has been decoded, so the whole read goes up",
                             "records_equal_resident_path": bool(rec_r.tobytes() == rec0.tobytes())},
               "gpu_launches": int(launches), "clocks": clocks, "paf_digest": digest(lines)}
        # the integer stages that dominate the map step, against the HBM roofline: algorithmic bytes of the stage / its time in the step run
        # alone (host-observed between synchronisations: includes the launch gaps of the stage's kernels)
        try:
            hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)); psrc = "measured (MEASURED_PEAKS.json)"
        except Exception:
            hbm_peak, psrc = 6650.0, "fallback"
        n_anc = float(stc["anchors"])
        out["roofline_map"] = {
            "anchor_fill_and_sort": {"kernels": "k_fill_anchors + k_sort_buckets<*>", "bound": "hbm", "unit": "GB/s", "peak": hbm_peak, "peak_source": psrc,
                                     "algorithmic_bytes": "per anchor: 8 B position gathered from the index + 8 B key written (fill); 8 B read + 8 B written (sort) = 32 B",
                                     "achieved": 32.0 * n_anc / (st_alone["sort_ms"] * 1e-3) / 1e9, "frac": 32.0 * n_anc / (st_alone["sort_ms"] * 1e-3) / 1e9 / hbm_peak,
                                     "traffic": "profiles/r02_map_kernels.md: k_sort_buckets<8192> 361 MB read + 301 MB written for 59 M keys (= algorithmic)"},
            "chaining": {"kernels": "k_chain_pre + k_chain_simple + k_chain", "bound": "hbm", "unit": "GB/s", "peak": hbm_peak, "peak_source": psrc,
                         "algorithmic_bytes": "per anchor: 8 B key read, 4 B score + 4 B predecessor + 1 B used flag written = 17 B",
                         "achieved": 17.0 * n_anc / (st_alone["chain_ms"] * 1e-3) / 1e9, "frac": 17.0 * n_anc / (st_alone["chain_ms"] * 1e-3) / 1e9 / hbm_peak,
                         "traffic": "profiles/r02_map_kernels.md: k_chain_pre 0.47 GB read + 0.52 GB written, k_chain 1.24 GB per launch (keys re-read by each of the three passes)",
                         "note": "latency / issue-bound integer recurrences (the chaining DP is sequential per read), not bandwidth-bound: the fraction is reported for "
                                 "completeness, the lever is the dependent-instruction chain per anchor (profiles/r02_map_kernels.md)"}}
        if not args.no_cpu_baseline:
            out["cpu_baseline"] = reference_map_sample(ds, cfg, raw, off, ln, lines, max(args.cpu_sample_reads, 64 * (os.cpu_count() or 1)), n_reads, steps=2)
    for c, x in pipes:
        x.close(); c.close()
    return out


def reference_map_sample(ds, cfg, raw, off, ln, gpu_lines, nsub, n_reads, steps=1):
    """The reference's own map loop -- kt_for(n_threads, map_worker_for, ...) through oracle/_ref/libra_refmap.so -- on the
    first nsub reads of the batch with every host thread; its PAF lines (minus mt:f:) are diffed against the GPU's."""
    import re
    from oracle import RefMap
    n_thr = os.cpu_count() or 1
    nsub = min(nsub, n_reads)
    if not RefMap.available():
        return {"value": None, "unit": "reads/s", "cores": n_thr, "kind": "reference", "sample": "unavailable: oracle/_ref/libra_refmap.so was not built"}
    rm = RefMap()
    idx = rm.index_load(ds.index_path)
    opt = rm.mapopt(**cfg["opt"])
    end = int((off[:nsub].astype(np.int64) + ln[:nsub].astype(np.int64)).max())      # reads need not be laid out in order (C4 mixes sources)
    sig = ds.decode(raw[:end])
    rm.map_reads(idx, opt, sig, off[:min(nsub, 32)], ln[:min(nsub, 32)], n_threads=n_thr)      # warm-up
    best = None
    for _ in range(steps):
        t0 = time.perf_counter()
        paf, mapped, chunks = rm.map_reads(idx, opt, sig, off[:nsub], ln[:nsub], n_threads=n_thr)
        t = time.perf_counter() - t0
        best = t if best is None else min(best, t)
    n1 = max(1, nsub // n_thr)
    t0 = time.perf_counter(); rm.map_reads(idx, opt, sig, off[:n1], ln[:n1], n_threads=1); t1 = time.perf_counter() - t0
    strip = lambda l: re.sub(r"mt:f:[0-9.eE+-]+\t", "", l)
    same = [strip(a) == b for a, b in zip(paf, gpu_lines[:nsub])]
    rm.index_free(idx)
    return {"value": nsub / best, "unit": "reads/s", "cores": n_thr, "kind": "reference",
            "sample": f"first {nsub} reads of the batch through the reference's own map loop (kt_for + map_worker_for, rmap.cpp:667-822, {n_thr} threads): "
                      "event detection, seeding, chaining and DTW as the reference does them; signal already decoded to float",
            "seconds": best, "one_thread_reads_per_s": n1 / t1, "thread_scaling": (nsub / best) / (n1 / t1),
            "mapped_frac": float(mapped.mean()), "mean_chunks": float(chunks.mean()),
            "paf_lines_equal_gpu": int(sum(same)), "paf_lines_compared": len(same), "parity_bit_exact": bool(all(same))}


def run_reference_map_arm(args):
    """--impl reference for the map-loop configs: the reference's map loop on all host threads, each step a bounded sample."""
    from workload.mapbench import CONFIGS, Dataset
    name = args.config
    cfg = CONFIGS[name]
    ds = Dataset(name, log=log)
    n_thr = os.cpu_count() or 1
    nsub = int(os.environ.get("RA_REF_SAMPLE_READS", 64 * n_thr))
    raw, off, ln, _, _ = ds.reads(nsub, seed=2)
    from oracle import RefMap
    if not RefMap.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libra_refmap.so (the reference's compiled map loop) was not built in this checkout"}), flush=True)
        return 0
    rm = RefMap(); idx = rm.index_load(ds.index_path); opt = rm.mapopt(**cfg["opt"])
    sig = ds.decode(raw)
    for _ in range(max(1, min(args.warmup, 2))):
        rm.map_reads(idx, opt, sig, off[:min(nsub, 4 * n_thr)], ln[:min(nsub, 4 * n_thr)], n_threads=n_thr)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        paf, mapped, chunks = rm.map_reads(idx, opt, sig, off, ln, n_threads=n_thr)
    dt = time.perf_counter() - t0
    n1 = max(1, nsub // n_thr)
    t1 = time.perf_counter(); rm.map_reads(idx, opt, sig, off[:n1], ln[:n1], n_threads=1); t1 = time.perf_counter() - t1
    val = nsub * args.steps / dt
    out = {"impl": "reference", "metric": "reads_per_s", "value": val, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": cfg["desc"] + "; the whole map loop per read batch", "reads_per_step": nsub, "reference_bases": int(sum(cfg["ref"])),
                      "n_sequences": len(cfg["ref"]), "preset": cfg["preset"], "note": cfg.get("note", "")},
           "cpu_baseline": {"value": val, "unit": "reads/s", "cores": n_thr, "kind": "reference",
                            "sample": f"{nsub} reads per step (same generator as the GPU arm) through the reference's own map loop: kt_for(n_threads, map_worker_for) "
                                      f"(rmap.cpp:916, :667-822) on {n_thr} threads, index loaded by the reference's ri_idx_load; signal already decoded to float",
                            "one_thread_reads_per_s": n1 / t1, "thread_scaling": val / (n1 / t1), "mapped_frac": float(mapped.mean()), "mean_chunks": float(chunks.mean())},
           "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)
    return 0


def _dev_copy(torch, dev, dst_ptr, src_ptr, n_floats):
    """device-to-device copy between two raw float arenas (the contexts' reference mirrors)"""
    class _V:
        def __init__(self, p):
            self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (p, False), "version": 2}
    torch.as_tensor(_V(dst_ptr), device=dev).copy_(torch.as_tensor(_V(src_ptr), device=dev))
    torch.cuda.synchronize(dev)


def run_reference_arm(args, config):
    """The reference's own CPU implementation of the path on all host threads; each step is a bounded sample of the same
    workload (same generators, same shapes) run inside one C call."""
    be, kind = cpu_backend()
    n_threads = os.cpu_count() or 1
    nsub = max(n_threads, min(args.reads_per_gpu, int(os.environ.get("RA_REF_SAMPLE_READS", 16 * n_threads))))
    ref_seq, ref_arena = build_reference()
    reads, coff, clen, crid = build_reads(ref_seq, nsub, seed=2)
    # events of the sample are needed to place the chains (same generator as the GPU arm): the oracle's peak list gives the
    # first sample of every event (event k starts at peaks[k-1])
    from oracle import Oracle
    orc = Oracle()
    ev, ev_off = orc.detect_events_batch(reads.sig, coff, clen)
    ev_pos = np.zeros(len(ev), np.uint32)
    for c in range(len(coff)):
        _, peaks = orc.detect_events(reads.sig[int(coff[c]):int(coff[c]) + int(clen[c])], return_peaks=True)
        n = int(ev_off[c + 1] - ev_off[c])
        if n > 1:
            ev_pos[int(ev_off[c]) + 1:int(ev_off[c]) + n] = peaks[:n - 1]
    anchors, chains = build_chains(ev_off, ev_pos, coff, crid, reads, nsub, REF_BASES - 6, seed=3)
    ch_rm, rfirst, _ = read_major(chains, nsub)
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        t_cpu, n_ev, sc = cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_threads)
    dt = time.perf_counter() - t0
    t_ev, _, _ = cpu_step(be, kind, reads, nsub, anchors, ch_rm, rfirst, ref_arena, n_threads, events_only=True)
    t_1, _, _ = cpu_step(be, kind, reads, max(1, nsub // n_threads), anchors, ch_rm, rfirst, ref_arena, 1)
    cells_ref, cells_all = useful_cells(ch_rm, anchors, sc)
    val = nsub * args.steps / dt
    one = max(1, nsub // n_threads) / t_1
    out = {"impl": "reference", "metric": "reads_per_s", "value": val, "unit": "reads/s", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "config": dict(config, reads_per_step=nsub),
           "cpu_baseline": {"value": val, "unit": "reads/s", "cores": n_threads, "kind": kind,
                            "sample": f"{nsub} reads per step (same generator and shape as the GPU arm); the whole step is ONE C call: per read on a "
                                      f"dynamically scheduled thread, detect_events per chunk + the reference's align_chain over every chunk round "
                                      f"(budget pre-check, early termination, running threshold); {n_threads} OpenMP threads",
                            "events_only_seconds_per_step": t_ev, "one_thread_reads_per_s": one, "thread_scaling": val / one,
                            "dtw_cells_evaluated": cells_ref, "dtw_cells_all_parts": cells_all,
                            "dtw_gcups": cells_ref / max(dt / args.steps - t_ev, 1e-9) / 1e9},
           "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
