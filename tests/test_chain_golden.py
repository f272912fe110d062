"""Chain-level parity and raw-signal decode (f3) against fixtures generated from the
reference's own compiled map loop (oracle/gen_golden_chains.py -> refmap_shim.cpp -> the unmodified rmap.cpp / rsig.cpp).

CPU part: the C restatement (oracle/ra_oracle.c) reproduces the fixtures bit for bit; where oracle/_ref/libra_refmap.so is
present it is also diffed live on fresh random inputs.  GPU part (-m gpu): the CUDA path through the C ABI reproduces the
same fixtures bit for bit."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits
from oracle import ANCHOR_DTYPE, DtwOpt, RefMap

MODES = [(1, 1, 10), (1, 1, 15), (1, 0, 10), (0, 1, 10), (0, 1, 15), (0, 0, 10)]


@pytest.fixture(scope="module")
def cg():
    g = np.load(os.path.join(GOLDEN, "chain_cases.npz"))
    from rawalign_b200 import CHAIN_DTYPE
    d = {k: g[k] for k in g.files}
    d["anchors"] = np.ascontiguousarray(g["anchors"]).view(ANCHOR_DTYPE).reshape(-1)
    d["chains"] = np.ascontiguousarray(g["chains"]).view(CHAIN_DTYPE).reshape(-1)
    return d


def _oracle_eval(oracle, cg, oo, running, min_score=None, min_keep=None):
    """rmap.cpp:515-524 around ora_align_chain."""
    anchors, chains, read, ref = cg["anchors"], cg["chains"], cg["read"], cg["ref"]
    out = np.zeros(len(chains), np.float32)
    best, grp = 0.0, None
    for k, ch in enumerate(chains):
        if grp != ch["group"]:
            best, grp = 0.0, ch["group"]
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        ms = best if running else float(min_score[k])
        s = oracle.align_chain(a, read[int(ch["read_off"]):], ref, oo, min_score=ms, score_fma=1)
        out[k] = s
        if s >= np.float32(oo.dtw_min_score) and s > best:
            best = float(s)
    return out


@pytest.mark.parametrize("border,fill,frac", MODES)
def test_oracle_align_chain_vs_reference_golden(oracle, cg, border, fill, frac):
    tag = f"b{border}f{fill}r{frac}"
    oo = DtwOpt(border, fill, frac / 100.0, 0.4, 20.0)
    got = _oracle_eval(oracle, cg, oo, True)
    assert np.array_equal(bits(got), bits(cg[f"score_run_{tag}"]))
    assert np.array_equal(got >= np.float32(20.0), cg[f"keep_run_{tag}"].astype(bool))
    got = _oracle_eval(oracle, cg, oo, False, cg["min_score"])
    assert np.array_equal(bits(got), bits(cg[f"score_ms_{tag}"]))
    assert (cg[f"score_run_{tag}"] == np.float32(-1e10)).sum() > 0        # early termination is exercised


def test_oracle_negative_min_score_golden(oracle, cg):
    oo = DtwOpt(1, 1, 0.10, 0.7, -5.0)
    got = _oracle_eval(oracle, cg, oo, True)
    assert np.array_equal(bits(got), bits(cg["score_run_neg"]))


def cigar_restatement(oracle, a, rd, ref, border, bonus=np.float32(0.4)):
    """align_chain(cigar=true), rmap.cpp:218-235 (global) and :282-299 (sparse), on top of the oracle's DTW_global_tb,
    including the reference's quirks."""
    wi, wj, wd, c_tot, n_al = [], [], [], np.float32(0), 0
    if border == 1:
        parts = len(a) - 1
        for p in range(parts):
            s, e = a[parts - p], a[parts - p - 1]
            c, i2, j2, d2 = oracle.dtw_global_tb(rd[s["query_position"]:e["query_position"] + 1], ref[s["target_position"]:e["target_position"] + 1])
            wi.append(i2.astype(np.uint64) + int(s["query_position"])); wj.append(j2.astype(np.uint64) + int(s["target_position"])); wd.append(d2)
            c_tot = np.float32(c_tot + c); n_al += int(e["query_position"]) - int(s["query_position"]) + 1
    else:
        s, e = a[-1], a[0]
        c, i2, j2, d2 = oracle.dtw_global_tb(rd[s["query_position"]:e["query_position"] + 1], ref[s["target_position"]:e["target_position"] + 1])
        i2 = i2.astype(np.uint64); j2 = j2.astype(np.uint64)
        i2[-1] += len(i2) * int(s["query_position"]); j2[-1] += len(j2) * int(s["target_position"])
        wi, wj, wd = [i2], [j2], [d2]; c_tot = c; n_al = int(e["query_position"]) - int(s["query_position"]) + 1
    score = np.float32(np.float64(np.float32(n_al)) * np.float64(bonus) - np.float64(c_tot))      # fused, rmap.cpp:306 as built
    return np.concatenate(wi), np.concatenate(wj), np.concatenate(wd), np.float32(c_tot), score


@pytest.mark.parametrize("border", [1, 0])
def test_cigar_restatement_vs_reference_golden(oracle, cg, border):
    off = cg[f"cigar_off_b{border}"]
    for n, k in enumerate(cg["cigar_sel"]):
        ch = cg["chains"][k]
        a = cg["anchors"][int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        if border == 0:
            a = a[:30]
        wi, wj, wd, cost, score = cigar_restatement(oracle, a, cg["read"][int(ch["read_off"]):], cg["ref"], border)
        s0, e0 = int(off[n]), int(off[n + 1])
        assert np.array_equal(wi, cg[f"cigar_i_b{border}"][s0:e0]) and np.array_equal(wj, cg[f"cigar_j_b{border}"][s0:e0])
        assert np.array_equal(bits(wd), bits(cg[f"cigar_d_b{border}"][s0:e0]))
        assert bits(cost) == bits(cg[f"cigar_cost_b{border}"][n]) and bits(score) == bits(cg[f"cigar_score_b{border}"][n])


@pytest.fixture(scope="module")
def dg():
    g = np.load(os.path.join(GOLDEN, "decode_cases.npz"))
    return [{k: g[f"c{i}_{k}"] for k in ("raw", "raw_off", "raw_len", "l_sig", "sig", "dig", "ran", "offset")} for i in range(int(g["n_cases"]))]


def test_oracle_decode_vs_reference_golden(oracle, dg):
    """ora_decode_raw == ri_read_sig (rsig.cpp:213-222) on FAST5 files read by the reference's own reader."""
    for c in dg:
        pos = 0
        for r in range(len(c["raw_len"])):
            a = int(c["raw_off"][r]); n = int(c["raw_len"][r])
            scale = np.float32(c["ran"][r]) / np.float32(c["dig"][r])                   # rsig.cpp:215, fp32 division
            d = oracle.decode_raw(c["raw"][a:a + n], c["offset"][r], scale)
            assert len(d) == int(c["l_sig"][r])
            assert np.array_equal(bits(d), bits(c["sig"][pos:pos + len(d)]))
            pos += len(d)


# ---- live differential against the compiled reference (this container only) ----------------------------------------
needs_refmap = pytest.mark.skipif(not RefMap.available(), reason="oracle/_ref/libra_refmap.so not built (no /root/reference here)")


@needs_refmap
def test_oracle_align_chain_live_differential(oracle):
    from oracle.gen_golden_chains import chain_workload
    rm = RefMap()
    read, ref, anchors, chains = chain_workload(n_reads=10, seed=77, ref_len=30_000, chains_per_read=5, gap_p=0.1)
    n = 0
    for border, fill, frac, bonus, dms in [(1, 1, 0.1, 0.4, 20.0), (1, 1, 0.3, 0.55, 5.0), (1, 0, 0.1, 0.4, 20.0), (0, 1, 0.05, 0.4, 20.0), (0, 0, 0.1, 0.9, 0.0)]:
        opt = rm.mapopt(dtw_border_constraint=border, dtw_fill_method=fill, dtw_band_radius_frac=frac, dtw_match_bonus=bonus, dtw_min_score=dms)
        oo = DtwOpt(border, fill, frac, bonus, dms)
        for ch in chains:
            a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
            if border == 0 and fill == 0:
                a = a[:25]
            for ms in (-1e10, 0.0, 35.0, 300.0):
                s1 = rm.align_chain(a, read[int(ch["read_off"]):], ref, opt, ms)
                s2 = oracle.align_chain(a, read[int(ch["read_off"]):], ref, oo, ms, score_fma=1)
                assert bits(s1) == bits(s2), (border, fill, frac, ms, s1, s2)
                n += 1
    assert n > 500


@needs_refmap
def test_reference_hotpath_batch_equals_per_call(oracle):
    """bench.py's reference arm (refmap_hotpath_batch: one threaded C call) gives what per-call reference functions give."""
    from oracle import Reference, SHIM_CHAIN_DTYPE
    from workload import synth
    rm = RefMap(); opt = rm.mapopt()
    ref_seq = synth.random_reference(40_000, seed=1)
    rd = synth.simulate_reads(ref_seq, 6, seed=5, max_chunks=3)
    coff, clen, crid = synth.chunk_table(rd.read_off, rd.read_len, max_chunks=3)
    ev_want, off_want = Reference(True).detect_events_batch(rd.sig, coff, clen)
    tot, _, _, cnt, ev, base = rm.hotpath_batch(rd.sig, rd.read_off[:-1], rd.read_len, None, None, None, None, opt, n_threads=3, max_chunks=3, want_events=True)
    assert tot == len(ev_want)
    got = np.concatenate([ev[int(base[r]):int(base[r]) + int(cnt[r])] for r in range(6)])
    assert np.array_equal(bits(got), bits(ev_want))


# ---- GPU: the CUDA path against the same fixtures ------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("border,fill,frac", MODES)
def test_gpu_align_chains_vs_reference_golden(ctx, cg, border, fill, frac):
    import rawalign_b200 as ra
    tag = f"b{border}f{fill}r{frac}"
    opt = ra.ri_mapopt_init(); opt.dtw_border_constraint = border; opt.dtw_fill_method = fill; opt.dtw_band_radius_frac = frac / 100.0
    ctx.set_opt(opt)
    try:
        score, keep, issued, useful = ctx.align_chains_batch(cg["read"], cg["ref"], cg["anchors"], cg["chains"], min_score_mode=1)
        score2, _, _, _ = ctx.align_chains_batch(cg["read"], cg["ref"], cg["anchors"], cg["chains"], min_score_mode=0, min_score=cg["min_score"])
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    assert np.array_equal(bits(score), bits(cg[f"score_run_{tag}"]))
    assert np.array_equal(keep.astype(bool), cg[f"keep_run_{tag}"].astype(bool))
    assert np.array_equal(bits(score2), bits(cg[f"score_ms_{tag}"]))


@pytest.mark.gpu
def test_gpu_negative_min_score_golden(ctx, cg):
    import rawalign_b200 as ra
    opt = ra.ri_mapopt_init(); opt.dtw_min_score = -5.0; opt.dtw_match_bonus = 0.7
    ctx.set_opt(opt)
    try:
        score, keep, _, _ = ctx.align_chains_batch(cg["read"], cg["ref"], cg["anchors"], cg["chains"], min_score_mode=1)
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    assert np.array_equal(bits(score), bits(cg["score_run_neg"])) and np.array_equal(keep.astype(bool), cg["keep_run_neg"].astype(bool))


@pytest.mark.gpu
@pytest.mark.parametrize("border", [1, 0])
def test_gpu_cigar_vs_reference_golden(ctx, cg, border):
    import rawalign_b200 as ra
    from rawalign_b200 import CHAIN_DTYPE
    sel = cg["cigar_sel"]
    chains = cg["chains"][sel].copy()
    anchors = cg["anchors"]
    if border == 0:                                              # the fixture cut global chains to their first 30 anchors
        chains["n_anchors"] = np.minimum(chains["n_anchors"], 30)
    opt = ra.ri_mapopt_init(); opt.dtw_border_constraint = border; opt.dtw_fill_method = 0
    ctx.set_opt(opt)
    try:
        score, cost, pi, pj, pd, off = ctx.align_chains_cigar_batch(cg["read"], cg["ref"], anchors, chains)
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    assert np.array_equal(off.astype(np.int64), cg[f"cigar_off_b{border}"])
    assert np.array_equal(pi, cg[f"cigar_i_b{border}"]) and np.array_equal(pj, cg[f"cigar_j_b{border}"])
    assert np.array_equal(bits(pd), bits(cg[f"cigar_d_b{border}"]))
    assert np.array_equal(bits(cost), bits(cg[f"cigar_cost_b{border}"])) and np.array_equal(bits(score), bits(cg[f"cigar_score_b{border}"]))


@pytest.mark.gpu
def test_gpu_decode_vs_reference_golden(ctx, oracle, dg):
    """k_decode_raw (+ K1 on what it kept) against signals the reference's ri_read_sig decoded from FAST5."""
    for c in dg:
        scale = (c["ran"].astype(np.float32) / c["dig"].astype(np.float32)).astype(np.float32)
        ev, off, lsig, first = ctx.detect_events_raw_batch(c["raw"], c["raw_off"], c["raw_len"], c["offset"].astype(np.float32), scale, 4000, 30)
        assert np.array_equal(lsig, c["l_sig"])
        # events of the reference-decoded signal, chunked as rmap.cpp:684-689 does
        coff, clen = [], []
        pos = 0
        for r in range(len(c["l_sig"])):
            n = int(c["l_sig"][r]); cs = 0; k = 0
            while cs < n and k < 30:
                coff.append(pos + cs); clen.append(min(4000, n - cs)); cs += 4000; k += 1
            pos += n
        ev_o, off_o = oracle.detect_events_batch(c["sig"], np.array(coff, np.uint64), np.array(clen, np.uint32))
        assert np.array_equal(off, off_o) and np.array_equal(bits(ev), bits(ev_o))
