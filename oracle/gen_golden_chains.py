"""Generates tests/golden/chain_cases.npz and decode_cases.npz from the reference's own map-loop code
(oracle/_ref/libra_refmap.so = refmap_shim.cpp + the unmodified rmap.cpp / rsig.cpp / ...).

Run where /root/reference exists:   python -m oracle.gen_golden_chains
Pins rows a7 (align_chain), a11 (CIGAR glue incl. quirks A.4-1/2), a12 (evaluation loop) and f3 (ri_read_sig decode).
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os
import tempfile

import numpy as np

from oracle import ANCHOR_DTYPE, SHIM_CHAIN_DTYPE, RefMap, build_refmap

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
MODES = [(1, 1), (1, 0), (0, 1), (0, 0)]          # (dtw_border_constraint, dtw_fill_method)


def chain_workload(n_reads=24, seed=41, ref_len=60_000, chains_per_read=4, gap_p=0.2):
    """Reads with one true chain each (so scores pass dtw_min_score and the running threshold bites) plus decoys."""
    from workload import synth
    rng = np.random.default_rng(seed)
    n_ev = rng.integers(100, 1800, n_reads)
    ev_off = np.concatenate(([0], np.cumsum(n_ev)))
    read = rng.normal(0, 1, int(ev_off[-1])).astype(np.float32)
    ref = rng.normal(0, 1, ref_len).astype(np.float32)
    anchors, chains = synth.synth_chains(n_ev, ev_off[:-1], len(ref), seed=seed + 1, chains_per_read=chains_per_read, gap_p=gap_p)
    for g in np.unique(chains["group"]):
        sel = np.nonzero(chains["group"] == g)[0]
        ch = chains[sel[int(rng.integers(0, 2))]]                 # the true chain is first or second in evaluation order
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        rd = read[int(ch["read_off"]):]
        for p in range(len(a) - 1, 0, -1):
            qs, qe = int(a["query_position"][p]), int(a["query_position"][p - 1])
            ts, te = int(a["target_position"][p]), int(a["target_position"][p - 1])
            seg = rd[qs:qe + 1]
            ref[ts:te + 1] = np.interp(np.linspace(0, len(seg) - 1, te - ts + 1), np.arange(len(seg)), seg) + rng.normal(0, 0.05, te - ts + 1)
    return read, ref.astype(np.float32), anchors, chains


def gen_chains(rm: RefMap):
    read, ref, anchors, chains = chain_workload()
    out = dict(read=read, ref=ref, anchors=anchors.view(np.uint32).reshape(-1, 2), chains=chains.view(np.uint8).reshape(len(chains), -1))
    sc = chains.astype(SHIM_CHAIN_DTYPE)
    for border, fill in MODES:
        for frac in (0.10, 0.15):
            if fill == 0 and frac != 0.10:
                continue
            opt = rm.mapopt(dtw_border_constraint=border, dtw_fill_method=fill, dtw_band_radius_frac=frac)
            s, k = rm.evaluate_chains_batch(read, ref, anchors, sc, opt, running=True, n_threads=4)
            tag = f"b{border}f{fill}r{int(frac * 100)}"
            out[f"score_run_{tag}"] = s; out[f"keep_run_{tag}"] = k
            ms = np.linspace(-50, 400, len(chains)).astype(np.float32)
            s2, k2 = rm.evaluate_chains_batch(read, ref, anchors, sc, opt, running=False, min_score=ms, n_threads=4)
            out[f"score_ms_{tag}"] = s2
            out["min_score"] = ms
            print(tag, "kept", int(k.sum()), "dropped(running)", int((s == np.float32(-1e10)).sum()), "dropped(ms)", int((s2 == np.float32(-1e10)).sum()))
    # a negative dtw_min_score and a different bonus
    opt = rm.mapopt(dtw_min_score=-5.0, dtw_match_bonus=0.7)
    s, k = rm.evaluate_chains_batch(read, ref, anchors, sc, opt, running=True, n_threads=2)
    out["score_run_neg"] = s; out["keep_run_neg"] = k
    # CIGAR (cigar=true): sparse and global, full fill (the only combination the reference supports for global)
    sel = np.arange(0, len(chains), 5)
    out["cigar_sel"] = sel
    for border in (1, 0):
        opt = rm.mapopt(dtw_border_constraint=border, dtw_fill_method=0)
        S, Cst, PI, PJ, PD, off = [], [], [], [], [], [0]
        for k in sel:
            ch = chains[k]
            a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
            if border == 0:
                a = a[:30]                                       # keep the full matrices small
            s, c, pi, pj, pd = rm.align_chain_cigar(a, read[int(ch["read_off"]):], ref, opt)
            S.append(s); Cst.append(c); PI.append(pi); PJ.append(pj); PD.append(pd); off.append(off[-1] + len(pi))
        out[f"cigar_score_b{border}"] = np.array(S, np.float32); out[f"cigar_cost_b{border}"] = np.array(Cst, np.float32)
        out[f"cigar_i_b{border}"] = np.concatenate(PI); out[f"cigar_j_b{border}"] = np.concatenate(PJ)
        out[f"cigar_d_b{border}"] = np.concatenate(PD); out[f"cigar_off_b{border}"] = np.array(off, np.int64)
        print("cigar border", border, "paths", len(sel), "elements", off[-1])
    np.savez_compressed(os.path.join(OUT, "chain_cases.npz"), **out)


def gen_decode(rm: RefMap):
    """int16 DAC samples -> FAST5 (our writer) -> the reference's open_sig / ri_read_sig (rsig.cpp:113,186-229)."""
    rng = np.random.default_rng(9)
    cases = []
    for dig, ran, off in [(8192.0, 1467.61, 10.0), (8192.0, 1402.882, -3.0), (2048.0, 748.58, 6.0), (8192.0, 1467.61, 0.0)]:
        lens = [int(x) for x in rng.integers(1, 9000, 5)] + [1, 7, 4000, 8001]
        raws = []
        for n in lens:
            pa = np.clip(rng.normal(90, 25, n), -20, 260)       # a good share outside (30, 200) and some near the edges
            raw = np.rint(pa * dig / ran - off).astype(np.int16)
            edge = rng.random(n) < 0.05
            raw[edge] = np.rint(np.where(rng.random(int(edge.sum())) < 0.5, 30.0, 200.0) * dig / ran - off).astype(np.int16) + rng.integers(-2, 3, int(edge.sum())).astype(np.int16)
            raws.append(raw)
        raw_len = np.array([len(r) for r in raws], np.uint32)
        raw_off = np.concatenate(([0], np.cumsum(raw_len)))[:-1].astype(np.uint64)
        raw = np.concatenate(raws)
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "x.fast5")
            rm.write_fast5(p, raw, raw_off, raw_len, digitisation=dig, range_=ran, offset=off)
            sig, l_sig, fdig, fran, foff, names = rm.read_fast5(p, len(raw_len) + 2, int(raw_len.sum()) + 16)
        assert len(l_sig) == len(raw_len) and names[0] == "read_00000000"
        cases.append(dict(raw=raw, raw_off=raw_off, raw_len=raw_len, l_sig=l_sig, sig=sig[: int(l_sig.sum())], dig=fdig, ran=fran, offset=foff))
        print("decode: dig/ran/off", dig, ran, off, "kept", int(l_sig.sum()), "of", int(raw_len.sum()))
    out = {}
    for i, c in enumerate(cases):
        for k, v in c.items():
            out[f"c{i}_{k}"] = v
    out["n_cases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(OUT, "decode_cases.npz"), **out)


def main():
    if not build_refmap():
        raise SystemExit("reference tree not available: cannot generate fixtures")
    os.makedirs(OUT, exist_ok=True)
    rm = RefMap()
    gen_chains(rm)
    gen_decode(rm)


if __name__ == "__main__":
    main()
