"""Seeded synthetic workload for the hot path: random reference, pore-model
expected signal, simulated R9.4-like raw signal, chunking, and candidate chains.

Nothing here reads /root/reference.  The 6-mer pore model is synthetic with the moments of the
R9.4 180mV 450bps template model the reference ships (mean level 90.2 pA, sd 12.9; mean level_stdv
2.16): the hot path only needs signal of that shape.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from rawalign_b200._lib import ANCHOR_DTYPE, CHAIN_DTYPE

K = 6
CHUNK_SIZE = 4000          # roptions.c:11
MAX_NUM_CHUNK = 30         # roptions.c:24
SAMPLES_PER_BASE = 4000.0 / 450.0


DAC_DIGITISATION, DAC_RANGE, DAC_OFFSET = 8192.0, 1467.61, 10.0     # R9.4 channel attributes


def pore_model(seed: int = 7):
    rng = np.random.default_rng(seed)
    mean = np.clip(rng.normal(90.2, 12.9, 4 ** K), 55.0, 130.0)
    stdv = np.clip(rng.normal(2.16, 0.5, 4 ** K), 1.0, 4.0)
    return mean.astype(np.float64), stdv.astype(np.float64)


def random_reference(n_bases: int, seed: int = 1) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, 4, n_bases, dtype=np.uint8)


def kmer_ranks(seq: np.ndarray) -> np.ndarray:
    n = len(seq) - K + 1
    r = np.zeros(n, np.int64)
    for i in range(K):
        r = r * 4 + seq[i:i + n]
    return r


def seq_to_sig(seq: np.ndarray, model_mean: np.ndarray, strand: int) -> np.ndarray:
    """Expected signal of a sequence, z-normalised per sequence in double (what ri_seq_to_sig,
    rsig.cpp:7-41, produces for the index): strand 0 walks the sequence as is, strand 1 walks the
    reverse complement."""
    s = seq if strand == 0 else (3 - seq[::-1])
    lv = model_mean[kmer_ranks(s)]
    m = lv.mean(); sd = lv.std()
    return ((lv - m) / sd).astype(np.float32)


@dataclass
class Reads:
    sig: np.ndarray          # float32 pA, reads concatenated (each read start 16-byte aligned)
    read_off: np.ndarray     # uint64 [n_reads+1] sample offsets (padded)
    read_len: np.ndarray     # uint32 samples per read
    ref_start: np.ndarray    # int64 first base on the sampled strand's coordinate
    strand: np.ndarray       # uint8
    n_bases: np.ndarray
    base_start: list         # per read: first sample of every base (int64), truth for anchor placement
    raw: np.ndarray = None       # int16 DAC samples, same layout as sig (what a FAST5 Signal dataset holds)
    dac_offset: np.ndarray = None  # float32 per read: ri_sig_t::offset
    dac_scale: np.ndarray = None   # float32 per read: range / digitisation (rsig.cpp:215)


def simulate_reads(ref: np.ndarray, n_reads: int, seed: int = 2, mean_len: float = 8000.0, max_chunks: int = MAX_NUM_CHUNK,
                   model=None) -> Reads:
    """Lognormal(mu=ln mean_len, sigma=0.5) read lengths in bases clipped to [1000, min(50000, |ref|)], uniform
    start, Bernoulli(0.5) strand, dwell 1+Poisson(7.9) samples per base, N(0, level_stdv) noise, clamp to
    (31,199) pA.  Only the first max_chunks*4000 samples are kept (the mapper never looks further)."""
    mm, ms = model or pore_model()
    rng = np.random.default_rng(seed)
    L = len(ref)
    nb = np.clip(rng.lognormal(np.log(mean_len), 0.5, n_reads), 1000, min(50000, L - K)).astype(np.int64)
    max_bases = int(max_chunks * CHUNK_SIZE / SAMPLES_PER_BASE * 1.3) + 64
    nb = np.minimum(nb, max_bases)
    strand = rng.integers(0, 2, n_reads).astype(np.uint8)
    start = (rng.random(n_reads) * (L - K + 1 - nb)).astype(np.int64)
    sigs, lens, bstarts = [], [], []
    for r in range(n_reads):
        a, b = int(start[r]), int(start[r] + nb[r]) + K - 1        # k-mer ranks of the read's bases only (the reference can be 250 Mb)
        ranks = kmer_ranks(ref[a:b] if strand[r] == 0 else (3 - ref[L - b:L - a][::-1]))
        dwell = 1 + rng.poisson(7.9, nb[r])
        lv = np.repeat(mm[ranks], dwell); sd = np.repeat(ms[ranks], dwell)
        s = lv + rng.normal(0.0, 1.0, len(lv)) * sd
        s = np.clip(s, 31.0, 199.0).astype(np.float32)[: max_chunks * CHUNK_SIZE]
        sigs.append(s); lens.append(len(s)); bstarts.append(np.concatenate(([0], np.cumsum(dwell)[:-1])))
    lens = np.array(lens, np.uint32)
    padded = (lens.astype(np.int64) + 3) // 4 * 4
    off = np.zeros(n_reads + 1, np.uint64)
    off[1:] = np.cumsum(padded)
    sig = np.zeros(int(off[-1]), np.float32)
    for r in range(n_reads):
        sig[int(off[r]):int(off[r]) + lens[r]] = sigs[r]
    # what the sequencer stores: raw = round(pA * digitisation / range - offset) as int16; the signal every
    # consumer sees is the reference's decode of it, pA = ((float)raw + offset) * (range / digitisation) in fp32
    dac_off = np.float32(DAC_OFFSET); dac_scale = np.float32(np.float32(DAC_RANGE) / np.float32(DAC_DIGITISATION))
    raw = np.rint(sig.astype(np.float64) * (DAC_DIGITISATION / DAC_RANGE) - DAC_OFFSET).astype(np.int16)
    dec = (raw.astype(np.float32) + dac_off) * dac_scale
    for r in range(n_reads):                                   # padding between reads stays out of the (30, 200) window check
        a, b = int(off[r]), int(off[r]) + int(lens[r])
        assert dec[a:b].min() > 30.0 and dec[a:b].max() < 200.0
    sig = dec.astype(np.float32)
    return Reads(sig, off, lens, start, strand, nb, bstarts, raw, np.full(n_reads, dac_off, np.float32), np.full(n_reads, dac_scale, np.float32))


def chunk_table(read_off: np.ndarray, read_len: np.ndarray, chunk_size: int = CHUNK_SIZE, max_chunks: int = MAX_NUM_CHUNK):
    """Chunks exactly as the reference's chunk loop cuts them (rmap.cpp:685-689): chunk c of a read is
    samples [c*chunk_size, min(l_sig, (c+1)*chunk_size)), at most max_chunks per read."""
    offs, lens, rid = [], [], []
    for r in range(len(read_len)):
        l = int(read_len[r]); base = int(read_off[r])
        nch = min(max_chunks, (l + chunk_size - 1) // chunk_size)
        for c in range(nch):
            offs.append(base + c * chunk_size); lens.append(min(chunk_size, l - c * chunk_size)); rid.append(r)
    return np.array(offs, np.uint64), np.array(lens, np.uint32), np.array(rid, np.uint32)


def synth_chains(n_read_events: np.ndarray, read_ev_off: np.ndarray, ref_len: int, seed: int = 3,
                 chains_per_read: int = 3, gap_p: float = 0.2, ref_per_event: float = 0.9):
    """Candidate chains with the shape sparse-border DTW sees: anchors strictly increasing in both
    coordinates (rmap.cpp:456-465), consecutive anchors 2+Geom(gap_p) read events apart, target gap =
    read gap scaled by ref_per_event plus jitter.  anchors[0] of a chain is its END (rmap.cpp:194-196).
    Each read gets chains_per_read chains at random reference positions (first one spans most of the
    read, the others are shorter decoys) so chain lengths vary the way chaining scores do."""
    rng = np.random.default_rng(seed)
    anchors, chains = [], []
    a_off = 0
    for r in range(len(n_read_events)):
        ne = int(n_read_events[r])
        if ne < 60:
            continue
        for c in range(chains_per_read):
            span = ne - 10 if c == 0 else int(rng.integers(30, max(31, ne // 2)))
            q0 = int(rng.integers(0, ne - span))
            gaps = 1 + rng.geometric(gap_p, size=span)          # >= 2
            q = q0 + np.concatenate(([0], np.cumsum(gaps)))
            q = q[q < q0 + span]
            if len(q) < 2:
                continue
            dq = np.diff(q)
            dt = np.maximum(1, np.rint(dq * ref_per_event + rng.normal(0, 0.7, len(dq)))).astype(np.int64)
            t_span = int(dt.sum())
            t0 = int(rng.integers(0, max(1, ref_len - t_span - 1)))
            t = t0 + np.concatenate(([0], np.cumsum(dt)))
            a = np.zeros(len(q), ANCHOR_DTYPE)
            a["query_position"] = q[::-1]; a["target_position"] = t[::-1]      # right-to-left, as traceback_chains emits
            ch = np.zeros(1, CHAIN_DTYPE)
            ch["anchor_off"] = a_off; ch["n_anchors"] = len(a); ch["group"] = r
            ch["read_off"] = read_ev_off[r]; ch["ref_off"] = 0
            anchors.append(a); chains.append(ch); a_off += len(a)
    if not chains:
        return np.zeros(0, ANCHOR_DTYPE), np.zeros(0, CHAIN_DTYPE)
    return np.concatenate(anchors), np.concatenate(chains)


def truth_chain(ev_sample_pos: np.ndarray, base_start: np.ndarray, ref_start: int, rng, gap_p: float = 0.2):
    """Anchors of the chain a perfect seeder would produce: read events q (2+Geom(gap_p) apart) paired with the
    reference k-mer position of the base being sequenced when event q starts.  Positions strictly increase in
    both coordinates (rmap.cpp:456-465); returned right-to-left like traceback_chains."""
    ne = len(ev_sample_pos)
    if ne < 12:
        return None
    gaps = 1 + rng.geometric(gap_p, size=ne)
    q = np.cumsum(gaps); q = q[q < ne - 1]
    if len(q) < 2:
        return None
    base = np.searchsorted(base_start, ev_sample_pos[q], "right") - 1
    t = ref_start + base
    keep = np.concatenate(([True], np.diff(t) > 0))
    # drop anchors whose target does not advance past the last kept one
    last = -1; sel = []
    for k in range(len(q)):
        if t[k] > last:
            sel.append(k); last = t[k]
    q = q[sel]; t = t[sel]
    if len(q) < 2:
        return None
    a = np.zeros(len(q), ANCHOR_DTYPE)
    a["query_position"] = q[::-1]; a["target_position"] = t[::-1]
    return a


# ------------------------------------------------------------------------------------------------------
# Files in the reference's own input formats (FASTA, k-mer model text), so its index builder and CLI can be driven
# ------------------------------------------------------------------------------------------------------
def write_fasta(path: str, seqs, names=None, width: int = 80):
    """seqs: list of uint8 arrays over {0,1,2,3} = ACGT."""
    lut = np.frombuffer(b"ACGT", np.uint8)
    with open(path, "wb") as f:
        for i, s in enumerate(seqs):
            f.write(b">" + (names[i] if names else f"ref{i}").encode() + b"\n")
            txt = lut[s]
            pad = (-len(txt)) % width
            rows = np.concatenate([txt, np.zeros(pad, np.uint8)]).reshape(-1, width)
            out = np.concatenate([rows, np.full((len(rows), 1), 10, np.uint8)], axis=1).reshape(-1)
            if pad:
                out = np.concatenate([out[: len(out) - pad - 1], [10]]).astype(np.uint8)
            f.write(out.tobytes())


def write_pore_model(path: str, model=None):
    """The k-mer model text format pore_model.cpp:57-82 parses: header line, then `kmer level_mean level_stdv sd_mean sd_stdv weight`."""
    mm, ms = model or pore_model()
    with open(path, "w") as f:
        f.write("kmer\tlevel_mean\tlevel_stdv\tsd_mean\tsd_stdv\tweight\n")
        for r in range(4 ** K):
            kmer = "".join("ACGT"[(r >> (2 * (K - 1 - i))) & 3] for i in range(K))
            f.write(f"{kmer}\t{mm[r]:.6f}\t{ms[r]:.6f}\t1.000000\t0.500000\t1000.0\n")


def pore_model_as_loaded(model=None) -> np.ndarray:
    """level_mean as the reference holds it after parsing the text written by write_pore_model: 6 decimals -> float."""
    mm, _ = model or pore_model()
    return np.array([float(f"{v:.6f}") for v in mm], np.float64).astype(np.float32)
