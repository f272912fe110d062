"""Datasets of BASELINE.json's configs for the map-loop benchmark and tests: synthetic reference FASTA, pore
model, simulated R9.4 reads as int16 DAC samples, and the index FILE in the reference's format.

The index is an input of the path (north_star: "the index format stays unchanged"): it is built by the reference's own indexer
(`oracle/_ref/rawalign -d`, the unmodified binary, when present) exactly as a user would build it, else written from the
pinned CPU restatement (oracle.build_flat_index + write_index_file).  Neither is inside any timed region.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import tempfile
import time

import numpy as np

from . import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RI_M_DTW_EVALUATE_CHAINS, RI_M_DTW_OUTPUT_CIGAR = 0x2, 0x4

# SURVEY.md 8(d) "Config -> flags" and "Synthetic inputs".  ref: list of sequence lengths; reads_from: which sequences reads are
# drawn from (weights); scale: documented reductions of the named size that keep a default run within minutes.
CONFIGS = {
    "C1": dict(desc="C1: ~30 kb viral-sized synthetic reference, -x viral --dtw-evaluate-chains, banded=0.1, sparse border",
               ref=[30_000], preset="viral", opt=dict(flag=RI_M_DTW_EVALUATE_CHAINS)),
    "C2": dict(desc="C2: 5 Mb synthetic reference, simulated R9.4 reads, -x sensitive --dtw-evaluate-chains, sparse border, banded=0.1",
               ref=[5_000_000], preset="sensitive", opt=dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_border_constraint=1, dtw_fill_method=1, dtw_band_radius_frac=0.10)),
    "C3": dict(desc="C3: C2 + --dtw-output-cigar", ref=[5_000_000], preset="sensitive",
               opt=dict(flag=RI_M_DTW_EVALUATE_CHAINS | RI_M_DTW_OUTPUT_CIGAR, dtw_border_constraint=1, dtw_fill_method=1, dtw_band_radius_frac=0.10)),
    "C4": dict(desc="C4: multi-genome synthetic index (20 genomes + one 30 kb viral target, contamination-style), -x fast, global border, full matrix",
               ref=[6_500_000, 1_625_000, 13_000_000, 3_250_000, 9_750_000, 2_275_000, 4_875_000, 8_125_000, 1_950_000, 2_925_000,
                    3_900_000, 2_600_000, 11_375_000, 3_575_000, 2_112_500, 5_850_000, 7_150_000, 1_787_500, 4_225_000, 3_087_500, 30_000],
               preset="fast", opt=dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_border_constraint=0, dtw_fill_method=0),
               viral_frac=0.10),
    "C5": dict(desc="C5: chromosome-scale synthetic reference, -x fast --dtw-evaluate-chains, banded=0.15",
               ref=[250_000_000], preset="fast", opt=dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_band_radius_frac=0.15)),
}


def _ref_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "rawalign")
    return p if os.path.exists(p) and os.access(p, os.X_OK) else None


class Dataset:
    def __init__(self, name: str, workdir: str | None = None, log=lambda *a: None):
        cfg = CONFIGS[name]
        self.name = name; self.cfg = cfg
        self.dir = workdir or os.path.join(tempfile.gettempdir(), f"ra_mapbench_{name}")
        os.makedirs(self.dir, exist_ok=True)
        t0 = time.time()
        self.seqs = [synth.random_reference(n, seed=1 + i) for i, n in enumerate(cfg["ref"])]
        self.fasta = os.path.join(self.dir, "ref.fa"); self.model = os.path.join(self.dir, "model.txt")
        self.index_path = os.path.join(self.dir, f"ref_{cfg['preset']}.ind")
        self.index_by = "cached"
        if not os.path.exists(self.index_path):
            synth.write_fasta(self.fasta, self.seqs); synth.write_pore_model(self.model)
            rb = _ref_binary()
            tmp_ind = self.index_path + f".{os.getpid()}.tmp"
            ok = False
            if rb:
                r = subprocess.run([rb, "-x", cfg["preset"], "-t", str(os.cpu_count() or 4), "-p", self.model, "-d", tmp_ind, self.fasta],
                                   capture_output=True, text=True)
                ok = r.returncode == 0 and os.path.exists(tmp_ind)
                if ok:
                    os.replace(tmp_ind, self.index_path); self.index_by = "the reference's indexer (oracle/_ref/rawalign -d)"
            if not ok:
                from oracle import build_flat_index, write_index_file
                write_index_file(tmp_ind, build_flat_index(self.seqs, synth.pore_model_as_loaded(), cfg["preset"]))
                os.replace(tmp_ind, self.index_path)
                self.index_by = "the pinned CPU restatement (oracle.build_flat_index)"
        log(f"[{name}] reference {sum(cfg['ref']) / 1e6:.1f} Mb in {len(self.seqs)} sequence(s); index {os.path.getsize(self.index_path) >> 20} MiB by {self.index_by} ({time.time() - t0:.1f}s)")

    def reads(self, n_reads: int, seed: int = 2):
        """n_reads simulated reads: int16 DAC samples, per-read offsets / lengths / channel attributes."""
        cfg = self.cfg
        vf = cfg.get("viral_frac", 0.0)
        if vf and len(self.seqs) > 1:
            rng = np.random.default_rng(seed + 1000)
            lens = np.array(cfg["ref"][:-1], np.float64)
            n_v = int(round(n_reads * vf))
            per = rng.multinomial(n_reads - n_v, lens / lens.sum())
            parts = [synth.simulate_reads(self.seqs[i], int(k), seed=seed * 100 + i) for i, k in enumerate(per) if k]
            if n_v:
                parts.append(synth.simulate_reads(self.seqs[-1], n_v, seed=seed * 100 + 99))
            raw = np.concatenate([p.raw for p in parts])
            off = np.concatenate([p.read_off[:-1] + np.uint64(sum(len(q.raw) for q in parts[:j])) for j, p in enumerate(parts)]).astype(np.uint64)
            ln = np.concatenate([p.read_len for p in parts]).astype(np.uint32)
            perm = rng.permutation(len(ln))
            off, ln = off[perm], ln[perm]
        else:
            rd = synth.simulate_reads(self.seqs[0], n_reads, seed=seed)
            raw, off, ln = rd.raw, rd.read_off[:-1].astype(np.uint64), rd.read_len.astype(np.uint32)
        n = len(ln)
        offset = np.full(n, np.float32(synth.DAC_OFFSET)); scale = np.full(n, np.float32(np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION)))
        return raw, off, ln, offset, scale

    @staticmethod
    def decode(raw):
        return ((raw.astype(np.float32) + np.float32(synth.DAC_OFFSET)) * (np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))).astype(np.float32)


def digest(lines) -> str:
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()[:16]
