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
    for _ in range(2):
        ctx.dtw_batch_dev(a_d.data_ptr(), b_d.data_ptr(), it_d.data_ptr(), items, c_d.data_ptr())
    ts = []
    for _ in range(5):
        ctx.dtw_batch_dev(a_d.data_ptr(), b_d.data_ptr(), it_d.data_ptr(), items, c_d.data_ptr())
        ts.append(ctx.last_timing()[2])
    ms = float(np.median(ts))
    g = cells / (ms * 1e-3) / 1e9
    out.append({"case": name, "items": n, "cells": int(cells), "ms": ms, "gcups": g, "frac_of_alu_roofline": g / ALU_PEAK})
    print(out[-1], flush=True)
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "microbench_dtw.json"), "w"), indent=1)
