"""Host-side mirror of the reference's interface for the hot path.

The reference is C/C++ (no Python); its boundary for this path is ``detect_events`` (revent.h:22),
the ``DTW_*`` functions (dtw.hpp:21-29) and ``align_chain`` (rmap.cpp:181).  ``Context`` exposes the
same names with the same argument meaning, each implemented by one call of the batched C ABI
(``include/rawalign_b200.h``) with a batch of one, plus the batch forms the re-batched map loop
uses.  Nothing here computes on the CPU: every result comes from the CUDA library.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import ANCHOR_DTYPE, CHAIN_DTYPE, ITEM_DTYPE, MapOpt, check

RI_M_DTW_BORDER_CONSTRAINT_GLOBAL = 0   # roptions.h:20-22
RI_M_DTW_BORDER_CONSTRAINT_SPARSE = 1
RI_M_DTW_BORDER_CONSTRAINT_LOCAL = 2
RI_M_DTW_FILL_METHOD_FULL = 0           # roptions.h:24-25
RI_M_DTW_FILL_METHOD_BANDED = 1


def ri_mapopt_init() -> MapOpt:
    """Defaults of ri_mapopt_init (roptions.c:5-61) for the hot-path fields."""
    o = MapOpt()
    _lib.load().ra_mapopt_init(C.byref(o))
    return o


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


def _f32(x) -> np.ndarray:
    return np.ascontiguousarray(x, dtype=np.float32)


def make_items(a_off, b_off, a_len, b_len, band_radius, exclude_last) -> np.ndarray:
    n = len(a_off)
    it = np.zeros(n, ITEM_DTYPE)
    it["a_off"] = a_off; it["b_off"] = b_off; it["a_len"] = a_len; it["b_len"] = b_len
    it["band_radius"] = band_radius; it["exclude_last"] = exclude_last
    return it


class Context:
    """One GPU context (``ra_ctx``): a stream, device arenas and the reference-signal mirror."""

    def __init__(self, device: int = 0, opt: Optional[MapOpt] = None):
        self._L = _lib.load()
        h = C.c_void_p()
        check(self._L.ra_init(int(device), C.byref(h)))
        self._h = h
        self.opt = opt or ri_mapopt_init()
        self.set_opt(self.opt)

    def close(self):
        if getattr(self, "_h", None):
            self._L.ra_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- options / reference mirror ------------------------------------------------------------
    def set_tuning(self, name: str, value: int):
        """ra_set_tuning: performance knobs that do not change results."""
        check(self._L.ra_set_tuning(self._h, name.encode(), int(value)))

    def set_opt(self, opt: MapOpt):
        check(self._L.ra_set_opt(self._h, C.byref(opt)))
        self.opt = opt

    @property
    def launch_count(self) -> int:
        return int(self._L.ra_launch_count(self._h))

    def last_timing(self):
        """(events_scan_ms, events_total_ms, dtw_ms) of the most recent batches, CUDA events on the library stream."""
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        check(self._L.ra_last_timing(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def ref_load_flat(self, ref: np.ndarray):
        ref = _f32(ref)
        check(self._L.ra_ref_load_flat(self._h, ref.ctypes.data, ref.size))

    def ref_load(self, fwd: Sequence[np.ndarray], rev: Sequence[np.ndarray]) -> np.ndarray:
        """Mirror ri_idx_t::forward_signals / reverse_signals; returns [n_seq, 2] arena offsets."""
        fwd = [_f32(x) for x in fwd]; rev = [_f32(x) for x in rev]
        n = len(fwd)
        fp = (C.c_void_p * n)(*[x.ctypes.data for x in fwd])
        rp = (C.c_void_p * n)(*[x.ctypes.data for x in rev])
        ln = np.array([len(x) for x in fwd], np.uint32)
        off = np.zeros(2 * n, np.uint64)
        check(self._L.ra_ref_load(self._h, n, C.cast(fp, C.c_void_p), C.cast(rp, C.c_void_p), ln.ctypes.data, off.ctypes.data))
        return off.reshape(n, 2)

    def ref_alloc(self, n_floats: int):
        check(self._L.ra_ref_alloc(self._h, int(n_floats)))

    def ref_device(self):
        p = C.c_void_p(); n = C.c_uint64()
        check(self._L.ra_ref_device(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    # ---- K1 --------------------------------------------------------------------------------------
    def detect_events_batch(self, sig, chunk_off, chunk_len, ev_out: Optional[np.ndarray] = None,
                            ev_off: Optional[np.ndarray] = None):
        """Batched detect_events.  Returns (events, ev_off) in CSR form."""
        sig = _f32(sig)
        chunk_off = np.ascontiguousarray(chunk_off, np.uint64)
        chunk_len = np.ascontiguousarray(chunk_len, np.uint32)
        n = len(chunk_off)
        if ev_out is None:
            ev_out = np.empty(max(int(chunk_len.sum()), 1), np.float32)
        if ev_off is None:
            ev_off = np.empty(n + 1, np.uint64)
        check(self._L.ra_detect_events_batch(self._h, sig.ctypes.data, sig.size, chunk_off.ctypes.data,
                                             chunk_len.ctypes.data, n, ev_out.ctypes.data, ev_out.size, ev_off.ctypes.data))
        return ev_out[:int(ev_off[n])], ev_off

    def detect_events_raw_batch(self, raw, raw_off, raw_len, offset, scale, chunk_size: int = 4000, max_chunks: int = 30,
                                ev_out: Optional[np.ndarray] = None):
        """Raw-signal form: int16 DAC samples in; ri_read_sig's conversion + (30,200) filter (rsig.cpp:213-222), the chunk
        loop of map_worker_for (rmap.cpp:684-689) and detect_events run on the device.
        Returns (events, ev_off, l_sig, chunk_first)."""
        raw = np.ascontiguousarray(raw, np.int16)
        raw_off = np.ascontiguousarray(raw_off, np.uint64); raw_len = np.ascontiguousarray(raw_len, np.uint32)
        offset = np.ascontiguousarray(offset, np.float32); scale = np.ascontiguousarray(scale, np.float32)
        n = len(raw_len)
        cap_chunks = int(np.minimum(max_chunks, (raw_len.astype(np.int64) + chunk_size - 1) // chunk_size).sum())
        if ev_out is None:
            ev_out = np.empty(max(int(raw_len.sum()) // 2 + cap_chunks + 16, 1), np.float32)
        ev_off = np.zeros(cap_chunks + 1, np.uint64)
        l_sig = np.zeros(n, np.uint32); chunk_first = np.zeros(n + 1, np.uint32)
        check(self._L.ra_detect_events_raw_batch(self._h, raw.ctypes.data, raw.size, raw_off.ctypes.data, raw_len.ctypes.data,
                                                 offset.ctypes.data, scale.ctypes.data, n, int(chunk_size), int(max_chunks),
                                                 l_sig.ctypes.data, chunk_first.ctypes.data, ev_out.ctypes.data, ev_out.size, ev_off.ctypes.data))
        nch = int(chunk_first[n])
        return ev_out[:int(ev_off[nch])], ev_off[:nch + 1], l_sig, chunk_first

    def detect_events(self, sig) -> np.ndarray:
        """detect_events(km, s_len, sig, opt, &n) for one chunk (revent.h:22): the event array
        (empty where the reference returns NULL)."""
        sig = _f32(sig)
        ev, _ = self.detect_events_batch(sig, np.array([0], np.uint64), np.array([len(sig)], np.uint32))
        return ev.copy()

    def detect_events_batch_dev(self, sig_ptr: int, chunk_off_ptr: int, chunk_len_ptr: int, n_chunks: int,
                                max_chunk_len: int, ev_out_ptr: int, ev_cap: int, ev_off_ptr: int) -> int:
        tot = C.c_uint64(0)
        check(self._L.ra_detect_events_batch_dev(self._h, sig_ptr, chunk_off_ptr, chunk_len_ptr, n_chunks, max_chunk_len,
                                                 ev_out_ptr, ev_cap, ev_off_ptr, C.byref(tot)))
        return int(tot.value)

    def last_event_positions(self, n_events: int) -> np.ndarray:
        """First sample (relative to its chunk) of every event of the last detect_events call."""
        pos = np.empty(max(int(n_events), 1), np.uint32)
        check(self._L.ra_last_event_positions(self._h, pos.ctypes.data, pos.size))
        return pos[:int(n_events)]

    # ---- K2 / K3 ---------------------------------------------------------------------------------
    def dtw_batch(self, a_arena, b_arena, items: np.ndarray, cost_out: Optional[np.ndarray] = None) -> np.ndarray:
        a_arena = _f32(a_arena)
        b_arena = None if b_arena is None else _f32(b_arena)
        items = np.ascontiguousarray(items, dtype=ITEM_DTYPE)
        if cost_out is None:
            cost_out = np.empty(len(items), np.float32)
        check(self._L.ra_dtw_batch(self._h, a_arena.ctypes.data, a_arena.size, _ptr(b_arena),
                                   0 if b_arena is None else b_arena.size, items.ctypes.data, len(items), cost_out.ctypes.data))
        return cost_out

    def dtw_batch_dev(self, a_ptr: int, b_ptr: Optional[int], items_ptr: int, items_host: np.ndarray, cost_ptr: int):
        items_host = np.ascontiguousarray(items_host, dtype=ITEM_DTYPE)
        check(self._L.ra_dtw_batch_dev(self._h, a_ptr, b_ptr, items_ptr, items_host.ctypes.data, len(items_host), cost_ptr))

    @staticmethod
    def dtw_cells(a_len: int, b_len: int, band_radius: int) -> int:
        return int(_lib.load().ra_dtw_cells(int(a_len), int(b_len), int(band_radius)))

    def _one(self, a, b, band_radius, exclude_last) -> np.float32:
        a, b = _f32(a), _f32(b)
        it = make_items([0], [0], [len(a)], [len(b)], [band_radius], [int(bool(exclude_last))])
        return self.dtw_batch(a, b, it)[0]

    def DTW_global(self, a_values, b_values, exclude_last_element: bool = False) -> np.float32:
        """dtw.hpp:21"""
        return self._one(a_values, b_values, -1, exclude_last_element)

    def DTW_global_slantedbanded_antidiagonalwise(self, a_values, b_values, band_radius: int,
                                                  exclude_last_element: bool = False) -> np.float32:
        """dtw.hpp:25"""
        if band_radius < 0:
            raise ValueError("band_radius must be >= 0 (dtw.cpp:277 asserts)")
        return self._one(a_values, b_values, int(band_radius), exclude_last_element)

    # ---- K4 --------------------------------------------------------------------------------------
    def dtw_tb_batch(self, a_arena, b_arena, items: np.ndarray, want_diff: bool = True):
        a_arena = _f32(a_arena)
        b_arena = None if b_arena is None else _f32(b_arena)
        items = np.ascontiguousarray(items, dtype=ITEM_DTYPE)
        cap = int((items["a_len"].astype(np.int64) + items["b_len"].astype(np.int64) - 1).sum()) if len(items) else 0
        cost = np.empty(len(items), np.float32)
        pi = np.empty(max(cap, 1), np.uint32); pj = np.empty(max(cap, 1), np.uint32)
        pd = np.empty(max(cap, 1), np.float32) if want_diff else None
        poff = np.zeros(len(items) + 1, np.uint64)
        check(self._L.ra_dtw_tb_batch(self._h, a_arena.ctypes.data, a_arena.size, _ptr(b_arena),
                                      0 if b_arena is None else b_arena.size, items.ctypes.data, len(items),
                                      cost.ctypes.data, pi.ctypes.data, pj.ctypes.data, _ptr(pd), cap, poff.ctypes.data))
        tot = int(poff[-1])
        return cost, pi[:tot], pj[:tot], (pd[:tot] if want_diff else None), poff

    def DTW_global_tb(self, a_values, b_values, exclude_last_element: bool = False):
        """dtw.hpp:28 -> (cost, i[], j[], difference[]) = dtw_result{cost, alignment}."""
        a, b = _f32(a_values), _f32(b_values)
        it = make_items([0], [0], [len(a)], [len(b)], [-1], [int(bool(exclude_last_element))])
        cost, pi, pj, pd, _ = self.dtw_tb_batch(a, b, it)
        return cost[0], pi.copy(), pj.copy(), pd.copy()

    # ---- chain level -----------------------------------------------------------------------------
    def align_chains_batch(self, read_arena, ref_arena, anchors: np.ndarray, chains: np.ndarray,
                           min_score_mode: int = 1, min_score: Optional[np.ndarray] = None, count_cells: bool = True):
        """Batched align_chain(cigar=false) + the evaluation loop of gen_chains (rmap.cpp:515-524).
        Returns (alignment_score, keep, cells_issued, cells_useful); the two cell counts are instrumentation
        (GCUPS numerator) and are 0 when count_cells is False."""
        read_arena = None if read_arena is None else _f32(read_arena)   # None: events resident from detect_events_batch
        ref_arena = None if ref_arena is None else _f32(ref_arena)
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        chains = np.ascontiguousarray(chains, dtype=CHAIN_DTYPE)
        ms = None if min_score is None else _f32(min_score)
        score = np.empty(len(chains), np.float32); keep = np.empty(len(chains), np.uint8)
        ci = C.c_uint64(0); cu = C.c_uint64(0)
        check(self._L.ra_align_chains_batch(self._h, _ptr(read_arena), 0 if read_arena is None else read_arena.size, _ptr(ref_arena),
                                            0 if ref_arena is None else ref_arena.size, anchors.ctypes.data, len(anchors),
                                            chains.ctypes.data, len(chains), int(min_score_mode), _ptr(ms),
                                            score.ctypes.data, keep.ctypes.data, C.byref(ci) if count_cells else None,
                                            C.byref(cu) if count_cells else None))
        return score, keep.astype(bool), int(ci.value), int(cu.value)

    def align_chains_batch_dev(self, read_ptr, ref_ptr, anchors_ptr: int, n_anchors: int, chains_ptr: int, n_chains: int,
                               group_first_ptr: int, n_groups: int, score_ptr: int, keep_ptr: Optional[int] = None, min_score_mode: int = 1,
                               min_score_ptr: Optional[int] = None, count_cells: bool = True):
        """Device-resident chain evaluation; returns (cells_issued, cells_useful) (0, 0 when count_cells is False)."""
        ci = C.c_uint64(0); cu = C.c_uint64(0)
        check(self._L.ra_align_chains_batch_dev(self._h, read_ptr, ref_ptr, anchors_ptr, int(n_anchors), chains_ptr, int(n_chains),
                                                group_first_ptr, int(n_groups),
                                                int(min_score_mode), min_score_ptr, score_ptr, keep_ptr,
                                                C.byref(ci) if count_cells else None, C.byref(cu) if count_cells else None))
        return int(ci.value), int(cu.value)

    def align_chain(self, anchors, read_events, ref_events, cigar: bool = False, min_score: float = -1e10):
        """align_chain(chain, ri, read_events, ..., opt, cigar, min_score) for one chain (rmap.cpp:181).
        anchors[0] is the chain end.  Returns alignment_score, or (alignment_score, cost, i, j, diff)
        when cigar is set."""
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        ch = np.zeros(1, CHAIN_DTYPE); ch["n_anchors"] = len(anchors)
        if not cigar:
            s, _, _, _ = self.align_chains_batch(read_events, ref_events, anchors, ch, 0, np.array([min_score], np.float32))
            return s[0]
        s, c, pi, pj, pd, _ = self.align_chains_cigar_batch(read_events, ref_events, anchors, ch)
        return s[0], c[0], pi, pj, pd

    def align_chains_cigar_batch(self, read_arena, ref_arena, anchors: np.ndarray, chains: np.ndarray):
        read_arena = _f32(read_arena)
        ref_arena = None if ref_arena is None else _f32(ref_arena)
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        chains = np.ascontiguousarray(chains, dtype=CHAIN_DTYPE)
        cap = 0
        for ch in chains:
            a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
            if len(a) == 0:
                continue
            q = a["query_position"].astype(np.int64); t = a["target_position"].astype(np.int64)
            cap += int((q[0] - q[-1]) + (t[0] - t[-1]) + 2 * len(a))
        score = np.empty(len(chains), np.float32); cost = np.empty(len(chains), np.float32)
        pi = np.empty(max(cap, 1), np.uint64); pj = np.empty(max(cap, 1), np.uint64); pd = np.empty(max(cap, 1), np.float32)
        poff = np.zeros(len(chains) + 1, np.uint64)
        check(self._L.ra_align_chains_cigar_batch(self._h, read_arena.ctypes.data, read_arena.size, _ptr(ref_arena),
                                                  0 if ref_arena is None else ref_arena.size, anchors.ctypes.data, len(anchors),
                                                  chains.ctypes.data, len(chains), score.ctypes.data, cost.ctypes.data,
                                                  pi.ctypes.data, pj.ctypes.data, pd.ctypes.data, cap, poff.ctypes.data))
        tot = int(poff[-1])
        return score, cost, pi[:tot], pj[:tot], pd[:tot], poff
