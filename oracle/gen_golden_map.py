"""Generates tests/golden/map_cases.npz from the reference's own compiled map loop (oracle/_ref/libra_refmap.so):
index tables (ri_idx_gen + ri_idx_dump), sketches (ri_sketch), PAF lines (map_worker_for) and per-chunk chain lists
(ri_map_frag / gen_chains) for a small two-sequence reference and a mix of clean, noisy and junk reads.

Run where /root/reference exists:   python -m oracle.gen_golden_map
Pins the map-loop restatement oracle/ra_oracle_map.c (rows f1, f2, f4 and the control flow between K1 and the DTW).
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

import os
import re
import tempfile

import numpy as np

from oracle import RI_M_DTW_EVALUATE_CHAINS, RefMap, build_refmap, parse_index_file

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SEQ_LENS = (60_000, 40_000)
CONFIGS = [("sensitive", dict(flag=RI_M_DTW_EVALUATE_CHAINS)),
           ("sensitive", dict(flag=0)),
           ("viral", dict(flag=RI_M_DTW_EVALUATE_CHAINS)),
           ("fast", dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_border_constraint=0, dtw_fill_method=0)),
           ("fast", dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_band_radius_frac=0.15)),
           ("sensitive", dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_border_constraint=0, dtw_fill_method=1))]


def map_workload():
    """Two reference sequences; 36 reads: clean ones from either sequence, noisy ones, junk (never map: 30 chunks), a very
    short one and one whose chunks fall under min_events."""
    from workload import synth
    rng = np.random.default_rng(77)
    seqs = [synth.random_reference(n, seed=11 + i) for i, n in enumerate(SEQ_LENS)]
    a = synth.simulate_reads(seqs[0], 20, seed=21, max_chunks=4)
    b = synth.simulate_reads(seqs[1], 12, seed=22, max_chunks=4)
    raws, lens = [], []
    for rd in (a, b):
        for r in range(len(rd.read_len)):
            o = int(rd.read_off[r]); raws.append(rd.raw[o:o + int(rd.read_len[r])].copy()); lens.append(int(rd.read_len[r]))
    dig, ran, off = synth.DAC_DIGITISATION, synth.DAC_RANGE, synth.DAC_OFFSET
    to_raw = lambda pa: np.rint(np.clip(pa, 31, 199) * dig / ran - off).astype(np.int16)
    to_pa = lambda raw: (raw.astype(np.float32) + np.float32(off)) * (np.float32(ran) / np.float32(dig))
    for r in (3, 9, 15, 25):                                      # noisy: several chunks before (if ever) mapping
        raws[r] = to_raw(to_pa(raws[r]) + rng.normal(0, 5.0, lens[r]))
    for n in (120_000, 30_000):                                   # junk: white level sequence, never maps
        lv = np.repeat(rng.normal(90, 13, n // 8 + 1), 8)[:n] + rng.normal(0, 2, n)
        raws.append(to_raw(lv)); lens.append(n)
    raws.append(raws[0][:37].copy()); lens.append(37)             # shorter than min_events worth of signal
    flat = to_raw(np.full(9000, 95.0) + rng.normal(0, 0.4, 9000)) # nearly constant: < min_events events per chunk
    raws.append(flat); lens.append(len(flat))
    lens = np.array(lens, np.uint32)
    pad = (lens.astype(np.int64) + 7) // 8 * 8
    raw_off = np.concatenate(([0], np.cumsum(pad)))[:-1].astype(np.uint64)
    raw = np.zeros(int(pad.sum()), np.int16)
    for r, x in enumerate(raws):
        raw[int(raw_off[r]):int(raw_off[r]) + len(x)] = x
    return seqs, raw, raw_off, lens


def table_checksum(ix) -> np.ndarray:
    """Order-sensitive digest of a flat index table: lengths plus position-weighted wrap-around sums."""
    def dig(a):
        a = np.ascontiguousarray(a, np.uint64)
        return np.array([len(a), np.bitwise_xor.reduce(a) if len(a) else 0,
                         (a * (np.arange(len(a), dtype=np.uint64) * np.uint64(2654435761) + np.uint64(1))).sum(dtype=np.uint64)], np.uint64)
    return np.concatenate([dig(ix["keys"]), dig(ix["key_off"]), dig(ix["vals"])])


def strip_mt(line: str) -> str:
    return re.sub(r"mt:f:[0-9.eE+-]+\t", "", line)


def main():
    from workload import synth
    if not build_refmap():
        raise SystemExit("reference tree not available: cannot generate fixtures")
    rm = RefMap()
    seqs, raw, raw_off, raw_len = map_workload()
    sig = ((raw.astype(np.float32) + np.float32(synth.DAC_OFFSET)) * (np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))).astype(np.float32)
    out = dict(seq0=seqs[0], seq1=seqs[1], raw=raw, raw_off=raw_off, raw_len=raw_len)
    with tempfile.TemporaryDirectory() as d:
        synth.write_fasta(d + "/ref.fa", seqs); synth.write_pore_model(d + "/model.txt")
        out["pore_vals"] = rm.pore_model(d + "/model.txt")
        lines = []
        for ci, (preset, kw) in enumerate(CONFIGS):
            idx = rm.index_build(d + "/ref.fa", d + "/model.txt", preset, dump=d + f"/ref_{preset}.ind")
            if f"keys_{preset}" not in out:
                ix = parse_index_file(d + f"/ref_{preset}.ind")
                out[f"keys_{preset}"] = table_checksum(ix)                   # full tables only for one preset (size)
                if preset == "sensitive":
                    out["keys_full"] = ix["keys"]; out["key_off_full"] = ix["key_off"]; out["vals_full"] = ix["vals"]
                    out["fwd1"] = ix["fwd"][1]; out["rev1"] = ix["rev"][1]
            opt = rm.mapopt(**kw)
            paf, mapped, chunks = rm.map_reads(idx, opt, sig, raw_off, raw_len, n_threads=4)
            lines.append("\n".join(strip_mt(l) for l in paf))
            print(preset, kw, "mapped", int(mapped.sum()), "of", len(mapped), "chunks", np.bincount(chunks).tolist())
            if ci == 0:                                            # per-chunk chain lists of a few reads (gen_chains output)
                CH, AN, NCH = [], [], []
                for r in (0, 3, 9, 25, 32):
                    a = int(raw_off[r])
                    per_chunk, _ = rm.read_chains(idx, opt, sig[a:a + int(raw_len[r])], max_chunks=6, stop_when_mapped=False)
                    for pc in per_chunk:
                        CH.append(pc["chains"]); AN.append(pc["anchors"]); NCH.append(len(pc["chains"]))
                out["chunk_chains"] = np.concatenate(CH).view(np.uint8).reshape(-1, CH[0].dtype.itemsize) if CH else np.zeros((0, 1), np.uint8)
                out["chunk_anchors"] = np.concatenate(AN).view(np.uint32).reshape(-1, 2)
                out["chunk_nchains"] = np.array(NCH, np.int32)
            rm.index_free(idx)
        out["paf"] = np.array(lines)
        # --dtw-output-cigar (rmap.cpp:715-717, :741-744): alns:f: and the aln:s: triples of chains[0] of every mapped read
        from oracle import RI_M_DTW_OUTPUT_CIGAR
        idx = rm.index_build(d + "/ref.fa", d + "/model.txt", "sensitive")
        paf, mapped, chunks = rm.map_reads(idx, rm.mapopt(flag=RI_M_DTW_EVALUATE_CHAINS | RI_M_DTW_OUTPUT_CIGAR), sig, raw_off, raw_len, n_threads=4)
        out["paf_cigar"] = np.array("\n".join(strip_mt(l) for l in paf))
        print("cigar lines", len(paf), "bytes", sum(len(l) for l in paf))
        rm.index_free(idx)
        # a sketch of real events (ri_sketch, the call of rmap.cpp:367)
        from oracle import Oracle
        ev = Oracle().detect_events(sig[:4000])
        x, y = rm.sketch(ev, e=6)
        out["sk_ev"] = ev; out["sk_x"] = x; out["sk_y"] = y
    np.savez_compressed(os.path.join(OUT, "map_cases.npz"), **out)
    print("map_cases.npz", os.path.getsize(os.path.join(OUT, "map_cases.npz")) >> 10, "KiB")


if __name__ == "__main__":
    main()
