"""Host-side mirror of the reference's map loop on top of the C ABI (ra_index_*, ra_map_batch).

``Index`` = ri_idx_t on the device (loaded from the reference's unchanged index file, or from a flat table);
``map_reads`` = kt_for(n_threads, map_worker_for, ...) over a batch of reads (rmap.cpp:916), returning the numbers of
every PAF line; ``paf_lines`` formats them exactly as rmap.cpp:730-758 / :954-972 do (minus the per-read wall time mt:f:).
Nothing here computes on the CPU beyond formatting.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import PAF_DTYPE, MapStats, RiMapOpt, check

RI_M_DTW_EVALUATE_CHAINS = 0x2     # roptions.h:12
RI_M_DTW_OUTPUT_CIGAR = 0x4


def ri_mapopt_full(**kw) -> RiMapOpt:
    """ri_mapopt_init (roptions.c:5-61), then keyword overrides."""
    o = RiMapOpt()
    _lib.load().ra_ri_mapopt_init(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


class Index:
    """ra_index: the seed table and the reference signals on the device, bound to one Context."""

    def __init__(self, ctx, handle, names=None):
        self._L = _lib.load(); self.ctx = ctx; self._h = handle
        info = (C.c_int32 * 9)()
        check(self._L.ra_index_info(self._h, info, None))
        self.w, self.e, self.n, self.q, self.lq, self.k, self.n_seq = [int(x) for x in info[:7]]
        self.seq_len = np.zeros(self.n_seq, np.uint32)
        check(self._L.ra_index_info(self._h, info, self.seq_len.ctypes.data))
        self.names = list(names) if names is not None else [f"ref{i}" for i in range(self.n_seq)]

    @classmethod
    def load_file(cls, ctx, path: str, names=None) -> "Index":
        """ri_idx_load (rawindex.cpp:317-377) straight to the device."""
        h = C.c_void_p()
        check(_lib.load().ra_index_load_file(ctx._h, path.encode(), C.byref(h)))
        if names is None:
            names = _names_from_index_file(path)
        return cls(ctx, h, names)

    @classmethod
    def load_flat(cls, ctx, tab: dict) -> "Index":
        fwd = [np.ascontiguousarray(a, np.float32) for a in tab["fwd"]]; rev = [np.ascontiguousarray(a, np.float32) for a in tab["rev"]]
        n_seq = len(fwd)
        fp = (C.c_void_p * n_seq)(*[a.ctypes.data for a in fwd]); rp = (C.c_void_p * n_seq)(*[a.ctypes.data for a in rev])
        keys = np.ascontiguousarray(tab["keys"], np.uint64); ko = np.ascontiguousarray(tab["key_off"], np.uint64); vals = np.ascontiguousarray(tab["vals"], np.uint64)
        sl = np.ascontiguousarray(tab["seq_len"], np.uint32)
        h = C.c_void_p()
        check(_lib.load().ra_index_load_flat(ctx._h, tab["w"], tab["e"], tab["n"], tab["q"], tab["lq"], tab["k"], n_seq, sl.ctypes.data,
                                             C.cast(fp, C.c_void_p), C.cast(rp, C.c_void_p), len(keys), keys.ctypes.data, ko.ctypes.data, vals.ctypes.data, C.byref(h)))
        return cls(ctx, h, tab.get("names"))

    @classmethod
    def build(cls, ctx, seqs, pore_vals, *, e: int, q: int, lq: int, k: int = 6, names=None) -> "Index":
        """ri_idx_gen on the device (ra_index_build): seqs = uint8 arrays of base codes 0..3 (ACGT), 4 = ambiguous; pore_vals =
        the 4^k level means of the pore model.  Parameters as ri_idx_init stores them (rawindex.cpp:9-24); w = n = 0."""
        ss = [np.ascontiguousarray(s, np.uint8) for s in seqs]
        n_seq = len(ss)
        sp = (C.c_void_p * n_seq)(*[a.ctypes.data for a in ss])
        sl = np.array([len(a) for a in ss], np.uint32)
        pv = np.ascontiguousarray(pore_vals, np.float32)
        if pv.size != 4 ** k:
            raise ValueError(f"pore model has {pv.size} levels, 4^{k} expected")
        h = C.c_void_p()
        check(_lib.load().ra_index_build(ctx._h, e, q, lq, k, n_seq, C.cast(sp, C.c_void_p), sl.ctypes.data, pv.ctypes.data, C.byref(h)))
        return cls(ctx, h, names)

    def clone(self, ctx) -> "Index":
        """ra_index_clone: the same index for another context, copied device to device."""
        h = C.c_void_p()
        check(self._L.ra_index_clone(self._h, ctx._h, C.byref(h)))
        return Index(ctx, h, self.names)

    def export(self) -> dict:
        """ra_index_export: the seed table as (keys ascending, key_off, positions ascending per key) -- parse_index_file's form."""
        nk = C.c_uint64(); nv = C.c_uint64()
        check(self._L.ra_index_export(self._h, C.byref(nk), C.byref(nv), None, None, None))
        keys = np.zeros(nk.value, np.uint64); ko = np.zeros(nk.value + 1, np.uint64); vals = np.zeros(max(nv.value, 1), np.uint64)
        check(self._L.ra_index_export(self._h, C.byref(nk), C.byref(nv), keys.ctypes.data, ko.ctypes.data, vals.ctypes.data))
        return dict(keys=keys, key_off=ko, vals=vals[:nv.value])

    def tables(self):
        """[(device pointer, bytes)] of the hash slots, the position array and the reference mirror (ra_index_tables)."""
        ptrs = (C.c_void_p * 3)(); nb = (C.c_uint64 * 3)()
        check(self._L.ra_index_tables(self._h, ptrs, nb))
        return [(int(ptrs[i] or 0), int(nb[i])) for i in range(3)]

    def counts(self):
        info = (C.c_int32 * 9)()
        check(self._L.ra_index_info(self._h, info, None))
        return [int(x) for x in info]

    @classmethod
    def alloc_like(cls, ctx, info, n_keys, n_vals, seq_len, names=None) -> "Index":
        """ra_index_alloc: an index of the same geometry whose tables are to be received by broadcast."""
        h = C.c_void_p()
        ia = (C.c_int32 * 9)(*info)
        sl = np.ascontiguousarray(seq_len, np.uint32)
        check(_lib.load().ra_index_alloc(ctx._h, ia, int(n_keys), int(n_vals), sl.ctypes.data, C.byref(h)))
        return cls(ctx, h, names)

    def close(self):
        if getattr(self, "_h", None):
            self._L.ra_index_destroy(self._h); self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def map_reads(self, opt: RiMapOpt, *, raw=None, raw_off=None, raw_len=None, offset=None, scale=None, sig=None, sig_off=None, sig_len=None,
                  resident_reads: int = 0, count_cells: bool = False):
        """One batch through ra_map_batch.  Either int16 DAC samples (raw, raw_off, raw_len, offset, scale) or float pA
        (sig, sig_off, sig_len), or resident_reads = n to map again the batch the previous call left on the device.
        Returns (records: PAF_DTYPE[n_reads], stats: dict)."""
        st = MapStats(); st.reserved = 1 if count_cells else 0
        if resident_reads:
            out = np.zeros(resident_reads, PAF_DTYPE)
            check(self._L.ra_map_batch(self.ctx._h, self._h, C.byref(opt), None, None, None, None, None, None, None, None,
                                       resident_reads, out.ctypes.data, C.byref(st)))
        elif raw is not None:
            raw = np.ascontiguousarray(raw, np.int16); ro = np.ascontiguousarray(raw_off, np.uint64); rl = np.ascontiguousarray(raw_len, np.uint32)
            of = np.ascontiguousarray(offset, np.float32); sc = np.ascontiguousarray(scale, np.float32)
            n = len(rl); out = np.zeros(n, PAF_DTYPE)
            check(self._L.ra_map_batch(self.ctx._h, self._h, C.byref(opt), raw.ctypes.data, ro.ctypes.data, rl.ctypes.data, of.ctypes.data, sc.ctypes.data,
                                       None, None, None, n, out.ctypes.data, C.byref(st)))
        else:
            sig = np.ascontiguousarray(sig, np.float32); so = np.ascontiguousarray(sig_off, np.uint64); sl = np.ascontiguousarray(sig_len, np.uint32)
            n = len(sl); out = np.zeros(n, PAF_DTYPE)
            check(self._L.ra_map_batch(self.ctx._h, self._h, C.byref(opt), None, None, None, None, None, sig.ctypes.data, so.ctypes.data, sl.ctypes.data,
                                       n, out.ctypes.data, C.byref(st)))
        return out, st.as_dict()

    def last_cigar(self, n_reads: int):
        """ra_map_last_cigar: (path_off[n_reads+1], path_i, path_j, path_diff) of the last map_reads call with RI_M_DTW_OUTPUT_CIGAR."""
        off = np.zeros(n_reads + 1, np.uint64)
        check(self._L.ra_map_last_cigar(self._h, n_reads, off.ctypes.data, None, None, None, 0))
        tot = int(off[-1])
        pi = np.zeros(max(tot, 1), np.uint64); pj = np.zeros(max(tot, 1), np.uint64); pd = np.zeros(max(tot, 1), np.float32)
        check(self._L.ra_map_last_cigar(self._h, n_reads, off.ctypes.data, pi.ctypes.data, pj.ctypes.data, pd.ctypes.data, tot))
        return off, pi[:tot], pj[:tot], pd[:tot]

    @staticmethod
    def aln_string(pi, pj, pd) -> str:
        """dtwresult_to_string (rmap.cpp:580-592): operator<< of size_t, size_t, float (6 significant digits)."""
        return "".join(f"({int(i)},{int(j)},{float(d):.6g})" for i, j, d in zip(pi, pj, pd))

    def paf_lines(self, rec: np.ndarray, read_names: Optional[Sequence[str]] = None, cigar=None):
        """The PAF lines of rmap.cpp:954-972 with the tags of :730-758 / :764-790, minus mt:f: (std::to_string(float) = %f)."""
        out = []
        for r, o in enumerate(rec):
            nm = read_names[r] if read_names is not None else f"read_{r}"
            if o["mapped"]:
                tags = (f"ci:i:{o['ci']}\tsl:i:{o['sl']}\tcm:i:{o['cm']}\tnc:i:{o['nc']}\ts1:f:{o['s1']:f}\ts2:f:{o['s2']:f}\tsm:f:{o['sm']:f}"
                        f"\tat:f:{o['at']:f}\taq:f:{o['aq']:f}")
                if cigar is not None:                                  # rmap.cpp:741-744
                    off, pi, pj, pd = cigar
                    a, b = int(off[r]), int(off[r + 1])
                    tags += f"\talns:f:{o['alns']:f}\taln:s:" + self.aln_string(pi[a:b], pj[a:b], pd[a:b])
                fs = int(o["fragment_start_position"]); fl = int(o["fragment_length"])
                rs = int(o["read_start_position"]); re_ = int(o["read_end_position"])
                out.append(f"{nm}\t{o['read_length']}\t{rs}\t{re_}\t{'-' if o['rev'] else '+'}\t{self.names[int(o['ref_id'])]}\t{self.seq_len[int(o['ref_id'])]}"
                           f"\t{fs}\t{(fs + fl) & 0xFFFFFFFF}\t{(re_ - rs - 1) & 0xFFFFFFFF}\t{fl}\t{o['mapq']}\t{tags}")
            elif o["nc"] >= 1:
                tags = (f"ci:i:{o['ci']}\tsl:i:{o['sl']}\tcm:i:{o['cm']}\tnc:i:{o['nc']}\ts1:f:{o['s1']:f}\ts2:f:{o['s2']:f}\tsm:f:{o['sm']:f}"
                        f"\tat:f:{o['at']:f}\taq:f:{o['aq']:f}")
                out.append(f"{nm}\t{o['read_length']}\t*\t*\t*\t*\t*\t*\t*\t*\t*\t{o['mapq']}\t{tags}")
            else:
                out.append(f"{nm}\t{o['read_length']}\t*\t*\t*\t*\t*\t*\t*\t*\t*\t{o['mapq']}\tci:i:{o['ci']}\tsl:i:{o['sl']}\tcm:i:0\tnc:i:0\ts1:f:0\ts2:f:0\tsm:f:0\tat:f:0\taq:f:0")
        return out


def _names_from_index_file(path: str):
    """Sequence names from the header of the reference's index file (rawindex.cpp:283-296)."""
    names = []
    with open(path, "rb") as f:
        assert f.read(2) == b"RI"
        pars = np.frombuffer(f.read(32), np.uint32)
        for _ in range(int(pars[6])):
            l = f.read(1)[0]
            names.append(f.read(l).decode())
            ln = int(np.frombuffer(f.read(4), np.uint32)[0])
            f.seek(8 * ln, 1)
    return names
