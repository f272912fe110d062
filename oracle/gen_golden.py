"""Generates tests/golden/*.npz from the reference's own objects (oracle/_ref/libra_ref*.so).

Run in a container that has /root/reference:   python -m oracle.gen_golden
The fixtures pin the oracle (and through it the CUDA path) on machines without the reference tree.
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os

import numpy as np

from oracle import ITEM_DTYPE, Reference, build_ref

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# check_dtw.cpp:183-237: the seven shape groups of run_various_random_tests
CHECKDTW_SHAPES = [(4, 4), (10, 10), (20, 10), (25, 10), (100, 100), (200, 50), (200, 30)]


def pack(arrs):
    """Concatenate variable-length arrays -> (flat, offsets)."""
    off = np.zeros(len(arrs) + 1, np.int64)
    off[1:] = np.cumsum([len(a) for a in arrs])
    flat = np.concatenate(arrs) if arrs else np.zeros(0, np.float32)
    return flat, off


def gen_dtw(ref: Reference):
    rng = np.random.default_rng(20240501)
    cases = []
    # (1) the reference unit test's generator: mt19937(seed) / mt19937(seed+1), check_dtw.cpp:139-141
    for gi, (al, bl) in enumerate(CHECKDTW_SHAPES):
        for s in range(12):
            seed = gi * 1428 + s
            a = ref.checkdtw_vector(al, seed); b = ref.checkdtw_vector(bl, seed + 1)
            for R in (1, 2, 3, 5, 8, 13, max(al, bl)):
                cases.append((a, b, R, s % 2))
    # (2) adversarial band shapes: every parity of R', m = 1, n = m, n >> m, tiny problems
    for _ in range(700):
        al = int(rng.integers(1, 400)); bl = int(rng.integers(1, 400)); R = int(rng.integers(0, 13))
        a = rng.uniform(-2.5, 2.5, al).astype(np.float32); b = rng.uniform(-2.5, 2.5, bl).astype(np.float32)
        cases.append((a, b, R, int(rng.integers(0, 2))))
    for al, bl in [(1, 1), (1, 7), (7, 1), (2, 2), (3, 4), (4, 3), (64, 64), (65, 33), (33, 65), (300, 2), (2, 300), (128, 127)]:
        for R in (0, 1, 2, 3, 4, 7):
            a = rng.uniform(-2.5, 2.5, al).astype(np.float32); b = rng.uniform(-2.5, 2.5, bl).astype(np.float32)
            cases.append((a, b, R, 0)); cases.append((a, b, R, 1))
    # (3) sparse-like tiny spans with R = 1
    for _ in range(400):
        al = int(2 + rng.geometric(0.2)); bl = int(2 + rng.geometric(0.2))
        a = rng.normal(0, 1, al).astype(np.float32); b = rng.normal(0, 1, bl).astype(np.float32)
        cases.append((a, b, 1, int(rng.integers(0, 2))))
    # (4) a few wide bands (radius = 0.1 * read length, as align_chain sets it)
    for al, bl in [(500, 550), (1500, 1300), (2600, 2500)]:
        a = rng.normal(0, 1, al).astype(np.float32); b = rng.normal(0, 1, bl).astype(np.float32)
        cases.append((a, b, max(1, int(al * 0.1)), 0))
    A, aoff = pack([c[0] for c in cases]); B, boff = pack([c[1] for c in cases])
    R = np.array([c[2] for c in cases], np.int32); ex = np.array([c[3] for c in cases], np.uint8)
    banded = np.array([ref.dtw_banded(c[0], c[1], c[2], bool(c[3])) for c in cases], np.float32)
    full = np.array([ref.dtw_global(c[0], c[1], bool(c[3])) for c in cases], np.float32)
    # traceback for the cases with < 2e4 cells
    tb_idx = np.array([k for k, c in enumerate(cases) if len(c[0]) * len(c[1]) <= 20000][::3], np.int64)
    tb_cost, tb_i, tb_j, tb_d = [], [], [], []
    for k in tb_idx:
        c = cases[k]
        cost, pi, pj, pd = ref.dtw_global_tb(c[0], c[1], bool(c[3]))
        tb_cost.append(cost); tb_i.append(pi); tb_j.append(pj); tb_d.append(pd)
    TI, tboff = pack(tb_i); TJ, _ = pack(tb_j); TD, _ = pack(tb_d)
    np.savez_compressed(os.path.join(OUT, "dtw_cases.npz"), a=A, a_off=aoff, b=B, b_off=boff, band_radius=R, exclude_last=ex,
                        banded_cost=banded, full_cost=full, tb_idx=tb_idx, tb_cost=np.array(tb_cost, np.float32),
                        tb_i=TI.astype(np.uint32), tb_j=TJ.astype(np.uint32), tb_diff=TD.astype(np.float32), tb_off=tboff)
    print("dtw cases:", len(cases), "tb cases:", len(tb_idx))


def sim_signal(n, rng):
    lv = rng.normal(90.2, 12.9, n // 5 + 2)
    dw = 1 + rng.poisson(7.9, len(lv))
    s = np.repeat(lv, dw)[:n]
    s = s + rng.normal(0, 2.16, len(s))
    return np.clip(s, 31, 199).astype(np.float32)


def gen_events(ref_fma: Reference, ref_nofma: Reference):
    rng = np.random.default_rng(20240502)
    chunks = []
    for _ in range(24):
        chunks.append(sim_signal(4000, rng))
    for n in (1, 2, 5, 6, 11, 12, 13, 16, 17, 31, 100, 257, 1023, 2048, 3999, 3001, 777):
        chunks.append(sim_signal(n + 40, rng)[:n])
    chunks.append(np.full(4000, 80.0, np.float32))                  # constant: variance -> FLT_MIN path
    chunks.append(np.full(50, 123.456, np.float32))
    step = np.where((np.arange(4000) // 40) % 2 == 0, 70.0, 110.0).astype(np.float32)   # noiseless steps: exact ties
    chunks.append(step)
    chunks.append((step + rng.normal(0, 0.01, 4000)).astype(np.float32))
    chunks.append(rng.uniform(31, 199, 4000).astype(np.float32))    # white noise: dense peaks
    chunks.append(np.clip(sim_signal(4000, rng) * 4.0, 31, 600).astype(np.float32))
    S, soff = pack(chunks)
    ev_f = [ref_fma.detect_events(c) for c in chunks]
    ev_n = [ref_nofma.detect_events(c) for c in chunks]
    EF, efoff = pack(ev_f); EN, enoff = pack(ev_n)
    np.savez_compressed(os.path.join(OUT, "event_cases.npz"), sig=S, sig_off=soff, ev_fma=EF, ev_fma_off=efoff,
                        ev_nofma=EN, ev_nofma_off=enoff)
    print("event chunks:", len(chunks), "events(fma):", len(EF), "chunks where fma != nofma:",
          sum(1 for x, y in zip(ev_f, ev_n) if x.tobytes() != y.tobytes()))


def main():
    if not build_ref():
        raise SystemExit("reference tree not available: cannot generate fixtures")
    os.makedirs(OUT, exist_ok=True)
    gen_dtw(Reference(True))
    gen_events(Reference(True), Reference(False))


if __name__ == "__main__":
    main()
