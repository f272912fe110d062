"""Kernel-level DTW micro-benchmark: GCUPS of ra_dtw_batch_dev on the reference's own
--performance-benchmark shapes (check_dtw.cpp:183-237) and a sparse-like stream.  Run on the GPU box.
Values uniform[-2.5, 2.5) (z-normalised scale), seed 42.  Cells = exact |cell set| (ra_dtw_cells)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rawalign_b200 as ra

dev = torch.device("cuda", 0)
ctx = ra.Context(0)
rng = np.random.default_rng(42)
ALU_PEAK = 148 * 128 * 1.965e9 / 4 / 1e9
out = []
shapes = [("banded 500x550 R=0.1", 500, 550, 50, 20000), ("banded 4000x4400 R=0.1", 4000, 4400, 400, 1000),
          ("banded 16000x16000 R=0.15", 16000, 16000, 2400, 64), ("full 500x550", 500, 550, -1, 4000),
          ("full 4000x4400", 4000, 4400, -1, 128)]
for name, al, bl, R, n in shapes:
    a = rng.uniform(-2.5, 2.5, (n, al)).astype(np.float32); b = rng.uniform(-2.5, 2.5, (n, bl)).astype(np.float32)
    items = ra.make_items(np.arange(n) * al, np.arange(n) * bl, np.full(n, al), np.full(n, bl), np.full(n, R), np.zeros(n))
    a_d = torch.from_numpy(a.reshape(-1)).to(dev); b_d = torch.from_numpy(b.reshape(-1)).to(dev)
    it_d = torch.from_numpy(items.view(np.uint8).reshape(-1)).to(dev); c_d = torch.empty(n, dtype=torch.float32, device=dev)
    cells = ra.Context.dtw_cells(al, bl, R) * n
    REPS = int(os.environ.get("RA_MB_REPS", 5))                    # RA_MB_REPS=1: one warm-up + one timed launch per shape (for ncu)
    for _ in range(2 if REPS > 1 else 1):
        ctx.dtw_batch_dev(a_d.data_ptr(), b_d.data_ptr(), it_d.data_ptr(), items, c_d.data_ptr())
    ts = []
    for _ in range(REPS):
        ctx.dtw_batch_dev(a_d.data_ptr(), b_d.data_ptr(), it_d.data_ptr(), items, c_d.data_ptr())
        ts.append(ctx.last_timing()[2])
    ms = float(np.median(ts))
    g = cells / (ms * 1e-3) / 1e9
    out.append({"case": name, "items": n, "cells": int(cells), "ms": ms, "gcups": g, "frac_of_alu_roofline": g / ALU_PEAK})
    print(out[-1], flush=True)
# (ii) sparse-like stream: anchor-to-anchor shapes, a_len, b_len ~ 2 + Geom(0.2), band radius 1, laid end to end
n = int(os.environ.get("RA_MB_SPARSE_ITEMS", 4_000_000))
al = (1 + rng.geometric(0.2, n)).astype(np.int64); bl = np.maximum(1, al + rng.integers(-1, 2, n)).astype(np.int64)
a_off = np.concatenate(([0], np.cumsum(al)[:-1])); b_off = np.concatenate(([0], np.cumsum(bl)[:-1]))
a = rng.uniform(-2.5, 2.5, int(al.sum())).astype(np.float32); b = rng.uniform(-2.5, 2.5, int(bl.sum())).astype(np.float32)
items = ra.make_items(a_off, b_off, al, bl, np.ones(n, int), np.ones(n, int))
a_d = torch.from_numpy(a).to(dev); b_d = torch.from_numpy(b).to(dev)
it_d = torch.from_numpy(items.view(np.uint8).reshape(-1)).to(dev); c_d = torch.empty(n, dtype=torch.float32, device=dev)
key = (al.astype(np.uint64) << np.uint64(32)) | bl.astype(np.uint64)
uniq, inv = np.unique(key, return_inverse=True)
cells = int(np.array([ra.Context.dtw_cells(int(k >> np.uint64(32)), int(k & np.uint64(0xffffffff)), 1) for k in uniq], np.int64)[inv].sum())
for hook, label in ((0, "sparse-like stream, R=1 (one item per thread)"), (4, "sparse-like stream, R=1 (generic warp-per-item kernel, test hook)")):
    opt = ra.ri_mapopt_init(); opt.reserved = hook
    c2 = ra.Context(0, opt)
    nn = n if hook == 0 else n // 8
    ts = []
    for rep in range(4 if int(os.environ.get("RA_MB_REPS", 5)) > 1 else 2):
        t0 = time.perf_counter()
        c2.dtw_batch_dev(a_d.data_ptr(), b_d.data_ptr(), it_d.data_ptr(), items[:nn], c_d.data_ptr())
        wall = (time.perf_counter() - t0) * 1e3
        if rep: ts.append((c2.last_timing()[2], wall))
    ms = float(np.median([t[0] for t in ts])); wall = float(np.median([t[1] for t in ts]))
    cc = cells if hook == 0 else int(cells * (nn / n))
    g = cc / (ms * 1e-3) / 1e9
    out.append({"case": label, "items": nn, "cells": cc, "ms": ms, "call_wall_ms_incl_host_sort": wall, "gcups": g, "frac_of_alu_roofline": g / ALU_PEAK})
    print(out[-1], flush=True)
    c2.close()
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "microbench_dtw.json"), "w"), indent=1)
