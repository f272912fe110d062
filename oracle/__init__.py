"""CPU oracle for RawAlign's hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package; the product package ``rawalign_b200`` never does.

``Oracle``   -- ctypes binding of ``oracle/libra_oracle.so`` (our C restatement, ``ra_oracle.c``).
``Reference``-- ctypes binding of ``oracle/_ref/libra_ref*.so`` (the reference's own ``dtw.cpp`` /
                ``revent.c`` compiled where they lie by ``oracle/Makefile``); present only when it was
                built in a container that has ``/root/reference``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")

ITEM_DTYPE = np.dtype([("a_off", "<u8"), ("b_off", "<u8"), ("a_len", "<u4"), ("b_len", "<u4"),
                       ("band_radius", "<i4"), ("exclude_last", "<u4")])
ANCHOR_DTYPE = np.dtype([("target_position", "<u4"), ("query_position", "<u4")])


class EvOpt(C.Structure):
    _fields_ = [("window_length1", C.c_uint32), ("window_length2", C.c_uint32),
                ("threshold1", C.c_float), ("threshold2", C.c_float), ("peak_height", C.c_float)]

    @classmethod
    def default(cls):  # roptions.c:49-53
        return cls(3, 6, 4.30265, 2.57058, 1.0)


class DtwOpt(C.Structure):
    _fields_ = [("dtw_border_constraint", C.c_uint32), ("dtw_fill_method", C.c_uint32),
                ("dtw_band_radius_frac", C.c_float), ("dtw_match_bonus", C.c_float),
                ("dtw_min_score", C.c_float)]

    @classmethod
    def default(cls):  # roptions.c:37-41
        return cls(1, 1, 0.10, 0.4, 20.0)


def build_oracle(force: bool = False) -> str:
    """Compile ra_oracle.c -> libra_oracle.so (gcc only; works on the GPU box too)."""
    so = os.path.join(_HERE, "libra_oracle.so")
    src = os.path.join(_HERE, "ra_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("ra_oracle.c", "ra_oracle_map.c", "ra_oracle.h", "ra_oracle_map.h")):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return so


def build_ref(ref_root: str = "/root/reference") -> bool:
    """Compile the reference's own sources into oracle/_ref/ when the tree is present."""
    if not os.path.isdir(os.path.join(ref_root, "src")):
        return os.path.exists(os.path.join(_HERE, "_ref", "libra_ref.so"))
    subprocess.check_call(["make", "-C", _HERE, "ref", f"REF={ref_root}"], stdout=subprocess.DEVNULL)
    return True


def _f32(x):
    return np.ascontiguousarray(x, dtype=np.float32)


class Oracle:
    """Our C restatement (oracle/ra_oracle.c)."""

    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.ora_detect_events.restype = C.c_uint32
        L.ora_detect_events.argtypes = [_f32p, C.c_uint32, C.POINTER(EvOpt), C.c_int, _f32p, _u32p, C.POINTER(C.c_uint32)]
        L.ora_detect_events_batch.restype = None
        L.ora_detect_events_batch.argtypes = [_f32p, _u64p, _u32p, C.c_uint32, C.POINTER(EvOpt), C.c_int, _f32p, _u64p]
        L.ora_dtw_global.restype = C.c_float
        L.ora_dtw_global.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int]
        L.ora_dtw_banded.restype = C.c_float
        L.ora_dtw_banded.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
        L.ora_dtw_cells.restype = C.c_uint64
        L.ora_dtw_cells.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.ora_dtw_global_tb.restype = C.c_uint32
        L.ora_dtw_global_tb.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int, C.POINTER(C.c_float), _u32p, _u32p, _f32p]
        L.ora_align_chain.restype = C.c_float
        L.ora_align_chain.argtypes = [C.c_void_p, C.c_uint32, _f32p, _f32p, C.POINTER(DtwOpt), C.c_float, C.c_int, C.POINTER(C.c_uint64)]
        L.ora_dtw_batch.restype = None
        L.ora_dtw_batch.argtypes = [_f32p, _f32p, C.c_void_p, C.c_uint32, _f32p]
        L.ora_decode_raw.argtypes = [C.c_void_p, C.c_uint32, C.c_float, C.c_float, _f32p]
        L.ora_decode_raw.restype = C.c_uint32

    def detect_events(self, sig, opt: EvOpt | None = None, contract: int = 1, return_peaks: bool = False):
        sig = _f32(sig)
        opt = opt or EvOpt.default()
        n = len(sig)
        ev = np.empty(max(n, 1), np.float32)
        peaks = np.empty(max(n, 1), np.uint32)
        npk = C.c_uint32(0)
        k = self.lib.ora_detect_events(sig, n, C.byref(opt), contract, ev, peaks, C.byref(npk))
        if return_peaks:
            return ev[:k].copy(), peaks[:npk.value].copy()
        return ev[:k].copy()

    def detect_events_batch(self, sig, chunk_off, chunk_len, opt: EvOpt | None = None, contract: int = 1):
        sig = _f32(sig)
        chunk_off = np.ascontiguousarray(chunk_off, np.uint64)
        chunk_len = np.ascontiguousarray(chunk_len, np.uint32)
        opt = opt or EvOpt.default()
        ev = np.empty(int(chunk_len.sum()) + 1, np.float32)
        off = np.empty(len(chunk_off) + 1, np.uint64)
        self.lib.ora_detect_events_batch(sig, chunk_off, chunk_len, len(chunk_off), C.byref(opt), contract, ev, off)
        return ev[:int(off[-1])].copy(), off

    def decode_raw(self, raw, offset: float, scale: float) -> np.ndarray:
        """ri_read_sig's conversion + (30,200) filter for one read (rsig.cpp:213-222)."""
        raw = np.ascontiguousarray(raw, np.int16)
        out = np.empty(len(raw) + 1, np.float32)
        n = self.lib.ora_decode_raw(raw.ctypes.data, len(raw), float(offset), float(scale), out)
        return out[:n].copy()

    def dtw_global(self, a, b, exclude_last=False):
        a, b = _f32(a), _f32(b)
        return np.float32(self.lib.ora_dtw_global(a, len(a), b, len(b), int(exclude_last)))

    def dtw_banded(self, a, b, band_radius, exclude_last=False, return_cells=False):
        a, b = _f32(a), _f32(b)
        cells = C.c_uint64(0)
        r = np.float32(self.lib.ora_dtw_banded(a, len(a), b, len(b), int(band_radius), int(exclude_last), C.byref(cells)))
        return (r, cells.value) if return_cells else r

    def dtw_cells(self, a_len, b_len, band_radius):
        return int(self.lib.ora_dtw_cells(int(a_len), int(b_len), int(band_radius)))

    def dtw_global_tb(self, a, b, exclude_last=False):
        a, b = _f32(a), _f32(b)
        cap = len(a) + len(b)
        pi = np.empty(cap, np.uint32); pj = np.empty(cap, np.uint32); pd = np.empty(cap, np.float32)
        cost = C.c_float(0)
        n = self.lib.ora_dtw_global_tb(a, len(a), b, len(b), int(exclude_last), C.byref(cost), pi, pj, pd)
        return np.float32(cost.value), pi[:n].copy(), pj[:n].copy(), pd[:n].copy()

    def align_chain(self, anchors, read, ref, opt: DtwOpt | None = None, min_score=-1e10, score_fma=0,
                    return_cells=False):
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        read, ref = _f32(read), _f32(ref)
        opt = opt or DtwOpt.default()
        cells = C.c_uint64(0)
        s = np.float32(self.lib.ora_align_chain(anchors.ctypes.data, len(anchors), read, ref, C.byref(opt),
                                                C.c_float(min_score), score_fma, C.byref(cells)))
        return (s, cells.value) if return_cells else s

    def dtw_batch(self, a_arena, b_arena, items):
        a_arena, b_arena = _f32(a_arena), _f32(b_arena)
        items = np.ascontiguousarray(items, dtype=ITEM_DTYPE)
        cost = np.empty(len(items), np.float32)
        self.lib.ora_dtw_batch(a_arena, b_arena, items.ctypes.data, len(items), cost)
        return cost


class Reference:
    """The reference's own objects (oracle/_ref/libra_ref.so), fma=True: reference flags on an FMA
    host; fma=False: same source with -ffp-contract=off."""

    @staticmethod
    def available(fma: bool = True) -> bool:
        return os.path.exists(os.path.join(_HERE, "_ref", "libra_ref.so" if fma else "libra_ref_nofma.so"))

    def __init__(self, fma: bool = True):
        path = os.path.join(_HERE, "_ref", "libra_ref.so" if fma else "libra_ref_nofma.so")
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_detect_events.restype = C.c_uint32
        L.ref_detect_events.argtypes = [_f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_float, _f32p]
        L.ref_detect_events_batch.restype = None
        L.ref_detect_events_batch.argtypes = [_f32p, _u64p, _u32p, C.c_uint32, _f32p, _u64p]
        for name in ("ref_dtw_global", "ref_dtw_global_slow"):
            f = getattr(L, name); f.restype = C.c_float
            f.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int]
        for name in ("ref_dtw_banded", "ref_dtw_slantedbanded"):
            f = getattr(L, name); f.restype = C.c_float
            f.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int, C.c_int]
        L.ref_dtw_global_tb.restype = C.c_uint32
        L.ref_dtw_global_tb.argtypes = [_f32p, C.c_uint32, _f32p, C.c_uint32, C.c_int, C.POINTER(C.c_float), _u32p, _u32p, _f32p]
        L.ref_dtw_batch.restype = None
        L.ref_dtw_batch.argtypes = [_f32p, _f32p, C.c_void_p, C.c_uint32, _f32p]
        L.ref_checkdtw_vector.restype = None
        L.ref_checkdtw_vector.argtypes = [C.c_int, C.c_uint, _f32p]

    def detect_events(self, sig, opt: EvOpt | None = None):
        sig = _f32(sig)
        opt = opt or EvOpt.default()
        ev = np.empty(max(len(sig), 1), np.float32)
        n = self.lib.ref_detect_events(sig, len(sig), opt.window_length1, opt.window_length2,
                                       opt.threshold1, opt.threshold2, opt.peak_height, ev)
        return ev[:n].copy()

    def detect_events_batch(self, sig, chunk_off, chunk_len):
        sig = _f32(sig)
        chunk_off = np.ascontiguousarray(chunk_off, np.uint64)
        chunk_len = np.ascontiguousarray(chunk_len, np.uint32)
        ev = np.empty(int(chunk_len.sum()) + 1, np.float32)
        off = np.empty(len(chunk_off) + 1, np.uint64)
        self.lib.ref_detect_events_batch(sig, chunk_off, chunk_len, len(chunk_off), ev, off)
        return ev[:int(off[-1])].copy(), off

    def dtw_global(self, a, b, exclude_last=False):
        a, b = _f32(a), _f32(b)
        return np.float32(self.lib.ref_dtw_global(a, len(a), b, len(b), int(exclude_last)))

    def dtw_banded(self, a, b, band_radius, exclude_last=False):
        a, b = _f32(a), _f32(b)
        return np.float32(self.lib.ref_dtw_banded(a, len(a), b, len(b), int(band_radius), int(exclude_last)))

    def dtw_global_tb(self, a, b, exclude_last=False):
        a, b = _f32(a), _f32(b)
        cap = len(a) + len(b)
        pi = np.empty(cap, np.uint32); pj = np.empty(cap, np.uint32); pd = np.empty(cap, np.float32)
        cost = C.c_float(0)
        n = self.lib.ref_dtw_global_tb(a, len(a), b, len(b), int(exclude_last), C.byref(cost), pi, pj, pd)
        return np.float32(cost.value), pi[:n].copy(), pj[:n].copy(), pd[:n].copy()

    def dtw_batch(self, a_arena, b_arena, items):
        a_arena, b_arena = _f32(a_arena), _f32(b_arena)
        items = np.ascontiguousarray(items, dtype=ITEM_DTYPE)
        cost = np.empty(len(items), np.float32)
        self.lib.ref_dtw_batch(a_arena, b_arena, items.ctypes.data, len(items), cost)
        return cost

    def checkdtw_vector(self, n, seed):
        out = np.empty(n, np.float32)
        self.lib.ref_checkdtw_vector(int(n), int(seed), out)
        return out


# ------------------------------------------------------------------------------------------------------
# The reference's whole map loop: oracle/_ref/libra_refmap.so (oracle/refmap_shim.cpp + every reference object)
# ------------------------------------------------------------------------------------------------------
class RiMapOpt(C.Structure):
    """ri_mapopt_t, roptions.h:39-87 (field for field; size checked against the compiled reference)."""
    _fields_ = [("bp_per_sec", C.c_uint32), ("sample_rate", C.c_uint32), ("chunk_size", C.c_uint32),
                ("min_events", C.c_uint32), ("max_gap_length", C.c_uint32), ("max_target_gap_length", C.c_uint32),
                ("chaining_band_length", C.c_uint32), ("max_num_skips", C.c_uint32), ("min_num_anchors", C.c_uint32),
                ("num_best_chains", C.c_uint32), ("min_chaining_score", C.c_float),
                ("step_size", C.c_uint32), ("max_num_chunk", C.c_uint32), ("min_chain_anchor", C.c_uint32),
                ("min_chain_anchor_out", C.c_uint32), ("dtw_border_constraint", C.c_uint32), ("dtw_fill_method", C.c_uint32),
                ("dtw_band_radius_frac", C.c_float), ("dtw_match_bonus", C.c_float), ("dtw_min_score", C.c_float),
                ("min_bestmap_ratio", C.c_float), ("min_bestmap_ratio_out", C.c_float),
                ("min_meanmap_ratio", C.c_float), ("min_meanmap_ratio_out", C.c_float),
                ("t_threshold", C.c_float), ("tn_samples", C.c_uint32), ("ttest_freq", C.c_uint32), ("tmin_reads", C.c_uint32),
                ("flag", C.c_int64), ("mini_batch_size", C.c_int64),
                ("window_length1", C.c_uint32), ("window_length2", C.c_uint32),
                ("threshold1", C.c_float), ("threshold2", C.c_float), ("peak_height", C.c_float)]


RI_M_DTW_EVALUATE_CHAINS = 0x2      # roptions.h:12-13
RI_M_DTW_OUTPUT_CIGAR = 0x4
SHIM_CHAIN_DTYPE = np.dtype([("anchor_off", "<u8"), ("n_anchors", "<u4"), ("group", "<u4"), ("read_off", "<u8"), ("ref_off", "<u8")])
CHAIN_OUT_DTYPE = np.dtype([("chaining_score", "<f4"), ("alignment_score", "<f4"), ("reference_sequence_index", "<u4"),
                            ("start_position", "<u4"), ("end_position", "<u4"), ("n_anchors", "<u4"), ("mapq", "<u4"),
                            ("strand", "<i4"), ("anchor_off", "<u8")])
# -x presets of main.cpp:131-150: (e, q, lq, w, n)
PRESETS = {"sensitive": (6, 9, 3, 0, 0), "fast": (7, 9, 3, 0, 0), "faster": (7, 9, 3, 5, 0), "viral": (5, 9, 3, 0, 0)}


class RefMap:
    """ctypes binding of oracle/_ref/libra_refmap.so: the reference's own align_chain / gen_chains / map_worker_for /
    ri_sketch / ri_idx_get / ri_seq_to_sig / ri_read_sig, compiled where they lie."""
    PATH = os.path.join(_HERE, "_ref", "libra_refmap.so")

    @staticmethod
    def available() -> bool:
        return os.path.exists(RefMap.PATH)

    def __init__(self):
        self.lib = L = C.CDLL(self.PATH)
        assert L.refmap_sizeof_mapopt() == C.sizeof(RiMapOpt), "RiMapOpt does not mirror ri_mapopt_t"
        vp = C.c_void_p
        L.refmap_index_build.restype = vp
        L.refmap_index_build.argtypes = [C.c_char_p, C.c_char_p] + [C.c_int] * 7 + [C.c_char_p]
        L.refmap_index_load.restype = vp; L.refmap_index_load.argtypes = [C.c_char_p]
        L.refmap_index_free.argtypes = [vp]
        L.refmap_index_info.argtypes = [vp, C.POINTER(C.c_int32)]
        L.refmap_index_seq.restype = C.c_uint32; L.refmap_index_seq.argtypes = [vp, C.c_uint32, C.c_char_p, C.c_uint32]
        L.refmap_index_signal.restype = C.POINTER(C.c_float); L.refmap_index_signal.argtypes = [vp, C.c_uint32, C.c_int]
        L.refmap_index_get.restype = C.c_int; L.refmap_index_get.argtypes = [vp, C.c_uint64, _u64p, C.c_int]
        L.refmap_pore_model.restype = C.c_int; L.refmap_pore_model.argtypes = [C.c_char_p, _f32p, C.c_uint32]
        L.refmap_sketch.restype = C.c_uint32
        L.refmap_sketch.argtypes = [_f32p, C.c_int] + [C.c_int] * 6 + [_u64p, _u64p, C.c_uint32]
        L.refmap_seq_to_sig.argtypes = [C.c_char_p, C.c_int, _f32p, C.c_int, C.c_int, C.POINTER(C.c_uint32), _f32p]
        L.refmap_align_chain.restype = C.c_float
        L.refmap_align_chain.argtypes = [vp, C.c_uint32, _f32p, _f32p, C.POINTER(RiMapOpt), C.c_int, C.c_float,
                                         C.POINTER(C.c_float), vp, vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
        L.refmap_evaluate_chains_batch.argtypes = [_f32p, _f32p, vp, vp, _u32p, C.c_uint32, C.POINTER(RiMapOpt), C.c_int, vp,
                                                   _f32p, _u8p, C.c_int]
        L.refmap_hotpath_batch.restype = C.c_uint64
        L.refmap_hotpath_batch.argtypes = [_f32p, _u64p, _u32p, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, vp, vp,
                                           C.POINTER(RiMapOpt), vp, vp, vp, vp, vp, C.c_int]
        L.refmap_reader_new.restype = vp; L.refmap_reader_new.argtypes = [vp, C.POINTER(RiMapOpt)]
        L.refmap_reader_chunk.restype = C.c_int; L.refmap_reader_chunk.argtypes = [vp, _f32p, C.c_uint32]
        L.refmap_reader_chains.restype = C.c_uint32
        L.refmap_reader_chains.argtypes = [vp, vp, C.c_uint32, vp, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.refmap_reader_events.restype = C.c_uint32; L.refmap_reader_events.argtypes = [vp, _f32p, C.c_uint32]
        L.refmap_reader_free.argtypes = [vp]
        L.refmap_map_reads.restype = C.c_uint64
        L.refmap_map_reads.argtypes = [vp, C.POINTER(RiMapOpt), _f32p, _u64p, _u32p, C.c_uint32, C.c_int, C.c_char_p, C.c_uint64, vp, vp]
        L.refmap_read_fast5.restype = C.c_int
        L.refmap_read_fast5.argtypes = [C.c_char_p, vp, C.c_uint64, _u32p, C.c_uint32, vp, vp, vp, vp, C.c_uint32]
        L.refmap_write_fast5.restype = C.c_int
        L.refmap_write_fast5.argtypes = [C.c_char_p, vp, _u64p, _u32p, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double]

    # ---- options ----
    def mapopt(self, **kw) -> RiMapOpt:
        o = RiMapOpt()
        self.lib.refmap_mapopt_init(C.byref(o))
        for k, v in kw.items():
            setattr(o, k, v)
        return o

    # ---- index ----
    def index_build(self, fasta: str, model: str, preset: str = "sensitive", k: int = 6, n_threads: int = 4, dump: str | None = None):
        e, q, lq, w, n = PRESETS[preset]
        h = self.lib.refmap_index_build(fasta.encode(), model.encode(), e, q, lq, w, n, k, n_threads, dump.encode() if dump else None)
        if not h:
            raise RuntimeError("ri_idx_reader_open failed")
        return h

    def index_load(self, path: str):
        h = self.lib.refmap_index_load(path.encode())
        if not h:
            raise RuntimeError("index load failed")
        return h

    def index_free(self, h):
        self.lib.refmap_index_free(h)

    def index_info(self, h) -> dict:
        a = (C.c_int32 * 9)()
        self.lib.refmap_index_info(h, a)
        return dict(zip(("b", "w", "e", "n", "q", "lq", "k", "flag", "n_seq"), list(a)))

    def index_signals(self, h):
        """[(name, len, forward_signals[i], reverse_signals[i])]: arrays of seq.len floats as ri_idx_dump writes them."""
        out = []
        for i in range(self.index_info(h)["n_seq"]):
            nm = C.create_string_buffer(256)
            ln = self.lib.refmap_index_seq(h, i, nm, 256)
            f = np.ctypeslib.as_array(self.lib.refmap_index_signal(h, i, 1), (ln,)).copy()
            r = np.ctypeslib.as_array(self.lib.refmap_index_signal(h, i, 0), (ln,)).copy()
            out.append((nm.value.decode(), ln, f, r))
        return out

    def index_get(self, h, hashval: int) -> np.ndarray:
        buf = np.empty(64, np.uint64)
        t = self.lib.refmap_index_get(h, int(hashval), buf, len(buf))
        if t > len(buf):
            buf = np.empty(t, np.uint64)
            self.lib.refmap_index_get(h, int(hashval), buf, len(buf))
        return buf[:t].copy()

    def pore_model(self, path: str) -> np.ndarray:
        lv = np.zeros(1 << 20, np.float32)
        k = self.lib.refmap_pore_model(path.encode(), lv, len(lv))
        return lv[: 1 << (2 * k)].copy()

    # ---- stages ----
    def sketch(self, ev, w=0, e=6, n=0, q=9, lq=3, k=6):
        ev = _f32(ev)
        x = np.empty(len(ev) * 3 + 8, np.uint64); y = np.empty_like(x)
        m = self.lib.refmap_sketch(ev, len(ev), w, e, n, q, lq, k, x, y, len(x))
        assert m <= len(x)
        return x[:m].copy(), y[:m].copy()

    def seq_to_sig(self, seq: bytes, pore_vals, k: int, strand: int) -> np.ndarray:
        out = np.zeros(len(seq) + 1, np.float32)
        n = C.c_uint32(0)
        self.lib.refmap_seq_to_sig(seq, len(seq), _f32(pore_vals), k, strand, C.byref(n), out)
        return out[: n.value].copy()

    # ---- align_chain ----
    def align_chain(self, anchors, read, ref, opt: RiMapOpt, min_score=-1e10):
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        return np.float32(self.lib.refmap_align_chain(anchors.ctypes.data, len(anchors), _f32(read), _f32(ref), C.byref(opt), 0,
                                                      C.c_float(min_score), None, None, None, None, 0, None))

    def align_chain_cigar(self, anchors, read, ref, opt: RiMapOpt):
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        q = anchors["query_position"].astype(np.int64); t = anchors["target_position"].astype(np.int64)
        cap = int((q[0] - q[-1]) + (t[0] - t[-1]) + 4 * len(anchors) + 16)
        pi = np.empty(cap, np.uint64); pj = np.empty(cap, np.uint64); pd = np.empty(cap, np.float32)
        n = C.c_uint64(0); cost = C.c_float(0)
        s = self.lib.refmap_align_chain(anchors.ctypes.data, len(anchors), _f32(read), _f32(ref), C.byref(opt), 1, C.c_float(-1e10),
                                        C.byref(cost), pi.ctypes.data, pj.ctypes.data, pd.ctypes.data, cap, C.byref(n))
        assert n.value <= cap
        return np.float32(s), np.float32(cost.value), pi[: n.value].copy(), pj[: n.value].copy(), pd[: n.value].copy()

    def evaluate_chains_batch(self, read_arena, ref_arena, anchors, chains, opt: RiMapOpt, running=True, min_score=None, n_threads=1):
        anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE)
        chains = np.ascontiguousarray(chains, dtype=SHIM_CHAIN_DTYPE)
        g = chains["group"]
        gfirst = np.concatenate((np.nonzero(np.concatenate(([True], g[1:] != g[:-1])))[0], [len(chains)])).astype(np.uint32) if len(chains) else np.zeros(1, np.uint32)
        score = np.empty(len(chains), np.float32); keep = np.empty(len(chains), np.uint8)
        ms = None if min_score is None else np.ascontiguousarray(min_score, np.float32)
        self.lib.refmap_evaluate_chains_batch(_f32(read_arena), _f32(ref_arena), anchors.ctypes.data, chains.ctypes.data, gfirst, len(gfirst) - 1,
                                              C.byref(opt), int(running), None if ms is None else ms.ctypes.data, score, keep, n_threads)
        return score, keep

    def hotpath_batch(self, sig, read_sig_off, read_l_sig, ref_arena, anchors, chains, round_first, opt: RiMapOpt, n_threads=1,
                      max_chunks=None, want_events=False):
        """One pass of the hot path over a batch of reads inside ONE C call (see refmap_shim.cpp)."""
        sig = _f32(sig)
        roff = np.ascontiguousarray(read_sig_off, np.uint64); rl = np.ascontiguousarray(read_l_sig, np.uint32)
        n_reads = len(rl)
        mc = int(max_chunks or opt.max_num_chunk)
        has = chains is not None and len(chains) > 0
        if has:
            anchors = np.ascontiguousarray(anchors, dtype=ANCHOR_DTYPE); chains = np.ascontiguousarray(chains, dtype=SHIM_CHAIN_DTYPE)
            round_first = np.ascontiguousarray(round_first, np.uint32)
            assert round_first.size == n_reads * (mc + 1)
            score = np.zeros(len(chains), np.float32); keep = np.zeros(len(chains), np.uint8)
            ref_arena = _f32(ref_arena)
        else:
            score = np.zeros(0, np.float32); keep = np.zeros(0, np.uint8)
        cnt = np.zeros(n_reads, np.uint32)
        ev = base = None
        if want_events:
            base = np.concatenate(([0], np.cumsum(np.minimum(rl.astype(np.int64), mc * int(opt.chunk_size)))))[:-1].astype(np.uint64)
            ev = np.zeros(int(np.minimum(rl.astype(np.int64), mc * int(opt.chunk_size)).sum()) + 1, np.float32)
        tot = self.lib.refmap_hotpath_batch(sig, roff, rl, n_reads, int(opt.chunk_size), mc,
                                            ref_arena.ctypes.data if has else None, anchors.ctypes.data if has else None,
                                            chains.ctypes.data if has else None, round_first.ctypes.data if has else None, C.byref(opt),
                                            score.ctypes.data if has else None, keep.ctypes.data if has else None,
                                            ev.ctypes.data if want_events else None, base.ctypes.data if want_events else None,
                                            cnt.ctypes.data, n_threads)
        return int(tot), score, keep, cnt, ev, base

    # ---- gen_chains / map loop ----
    def read_chains(self, idx, opt: RiMapOpt, sig, max_chunks=None, stop_when_mapped=True):
        """The chunk loop of map_worker_for (rmap.cpp:685-693) for one read; returns per chunk (chains, anchors, mapped)."""
        sig = _f32(sig)
        h = self.lib.refmap_reader_new(idx, C.byref(opt))
        out = []
        cs = int(opt.chunk_size); mc = int(max_chunks or opt.max_num_chunk)
        try:
            c = 0; start = 0
            while start < len(sig) and c < mc:
                end = min(len(sig), start + cs)
                mapped = self.lib.refmap_reader_chunk(h, sig[start:end], end - start)
                ch = np.zeros(64, CHAIN_OUT_DTYPE); an = np.zeros(1 << 16, ANCHOR_DTYPE)
                ne = C.c_uint32(0); ro = C.c_uint32(0)
                n = self.lib.refmap_reader_chains(h, ch.ctypes.data, len(ch), an.ctypes.data, len(an), C.byref(ne), C.byref(ro))
                tot = int(ch["n_anchors"][:n].sum())
                out.append(dict(chains=ch[:n].copy(), anchors=an[:tot].copy(), mapped=bool(mapped), n_events=ne.value, offset=ro.value))
                if mapped and stop_when_mapped:
                    break
                start += cs; c += 1
            ev = np.empty(max(1, out[-1]["n_events"] if out else 1), np.float32)
            n = self.lib.refmap_reader_events(h, ev, len(ev))
            return out, ev[:n].copy()
        finally:
            self.lib.refmap_reader_free(h)

    def map_reads(self, idx, opt: RiMapOpt, sig, sig_off, l_sig, n_threads=1):
        sig = _f32(sig); so = np.ascontiguousarray(sig_off, np.uint64); ls = np.ascontiguousarray(l_sig, np.uint32)
        cap = 4096 * len(ls) + (1 << 20)
        if opt.flag & RI_M_DTW_OUTPUT_CIGAR:
            cap += 64 * int(ls.sum()) // 4
        buf = C.create_string_buffer(cap)
        mapped = np.zeros(len(ls), np.uint8); chunks = np.zeros(len(ls), np.uint32)
        n = self.lib.refmap_map_reads(idx, C.byref(opt), sig, so, ls, len(ls), n_threads, buf, cap, mapped.ctypes.data, chunks.ctypes.data)
        return buf.raw[:n].decode().splitlines(), mapped, chunks

    # ---- FAST5 ----
    def write_fast5(self, path, raw, raw_off, raw_len, first_id=0, digitisation=8192.0, range_=1467.61, offset=10.0):
        raw = np.ascontiguousarray(raw, np.int16)
        rc = self.lib.refmap_write_fast5(path.encode(), raw.ctypes.data, np.ascontiguousarray(raw_off, np.uint64),
                                         np.ascontiguousarray(raw_len, np.uint32), len(raw_len), first_id, digitisation, range_, offset)
        if rc:
            raise RuntimeError("write_fast5 failed")

    def read_fast5(self, path, n_reads_cap, sig_cap):
        l_sig = np.zeros(n_reads_cap, np.uint32); sig = np.zeros(sig_cap, np.float32)
        dig = np.zeros(n_reads_cap, np.float32); ran = np.zeros(n_reads_cap, np.float32); off = np.zeros(n_reads_cap, np.float32)
        names = C.create_string_buffer(64 * n_reads_cap)
        n = self.lib.refmap_read_fast5(path.encode(), sig.ctypes.data, sig_cap, l_sig, n_reads_cap, dig.ctypes.data, ran.ctypes.data,
                                       off.ctypes.data, C.cast(names, C.c_void_p), 64)
        if n < 0:
            raise RuntimeError("open_sig failed")
        nm = [names.raw[i * 64:(i + 1) * 64].split(b"\0")[0].decode() for i in range(n)]
        return sig, l_sig[:n], dig[:n], ran[:n], off[:n], nm


def build_refmap(ref_root: str = "/root/reference") -> bool:
    """oracle/_ref/libra_refmap.so + oracle/_ref/rawalign (needs the reference tree; about two minutes the first time)."""
    if not os.path.isdir(os.path.join(ref_root, "src")):
        return RefMap.available()
    subprocess.check_call(["make", "-C", _HERE, "refmap", f"REF={ref_root}"], stdout=subprocess.DEVNULL)
    return True


# ------------------------------------------------------------------------------------------------------
# Map-loop restatement (oracle/ra_oracle_map.c): sketch, index lookup, chaining, primary chains, mapq, chunk loop
# ------------------------------------------------------------------------------------------------------
PAF_DTYPE = np.dtype([("mapped", "<u4"), ("ref_id", "<u4"), ("read_length", "<u4"), ("read_start_position", "<u4"),
                      ("read_end_position", "<u4"), ("fragment_start_position", "<u4"), ("fragment_length", "<u4"), ("mapq", "<u4"),
                      ("rev", "<u4"), ("ci", "<u4"), ("sl", "<u4"), ("cm", "<u4"), ("nc", "<u4"),
                      ("s1", "<f4"), ("s2", "<f4"), ("sm", "<f4"), ("at", "<f4"), ("aq", "<f4"), ("alns", "<f4"),
                      ("n_events", "<u4"), ("cells", "<u8")], align=True)


def parse_index_file(path: str) -> dict:
    """The reference's index file (ri_idx_dump, rawindex.cpp:275-315): magic "RI", 8 x u32 (w, e, n, q, lq, k, n_seq, flag),
    per sequence (u8 name length, name, u32 len, len fp32 forward_signals, len fp32 reverse_signals), then for each of the
    2^14 buckets: i32 n, n x u64 p[], u32 size, size x (u64 key, u64 val).  key = (hash >> 14) << 1 | single; val = the one
    position (single) or p_start << 32 | count.  Returns the flat table sorted by hash."""
    buf = np.fromfile(path, np.uint8)
    assert bytes(buf[:2]) == b"RI"
    pars = buf[2:34].view(np.uint32)
    w, e, n, q, lq, k, n_seq, flag = [int(x) for x in pars]
    o = 34
    names, lens, fwd, rev = [], [], [], []
    for _ in range(n_seq):
        l = int(buf[o]); o += 1
        names.append(bytes(buf[o:o + l]).decode()); o += l
        ln = int(buf[o:o + 4].view(np.uint32)[0]); o += 4
        fwd.append(buf[o:o + 4 * ln].view(np.float32).copy()); o += 4 * ln
        rev.append(buf[o:o + 4 * ln].view(np.float32).copy()); o += 4 * ln
        lens.append(ln)
    b = 14
    H, V, C = [], [], []
    for i in range(1 << b):
        nb = int(buf[o:o + 4].view(np.int32)[0]); o += 4
        p = buf[o:o + 8 * nb].view(np.uint64); o += 8 * nb
        size = int(buf[o:o + 4].view(np.uint32)[0]); o += 4
        if size == 0:
            continue
        kv = buf[o:o + 16 * size].view(np.uint64).reshape(size, 2); o += 16 * size
        key, val = kv[:, 0], kv[:, 1]
        h = ((key >> np.uint64(1)) << np.uint64(b)) | np.uint64(i)
        single = (key & np.uint64(1)).astype(bool)
        cnt = np.where(single, 1, (val & np.uint64(0xFFFFFFFF))).astype(np.int64)
        start = (val >> np.uint64(32)).astype(np.int64)
        vals = np.empty(int(cnt.sum()), np.uint64)
        offs = np.concatenate(([0], np.cumsum(cnt)))
        vals[offs[:-1][single]] = val[single]
        for j in np.nonzero(~single)[0]:
            vals[offs[j]:offs[j + 1]] = p[start[j]:start[j] + cnt[j]]
        H.append(h); C.append(cnt); V.append(vals)
    assert o == len(buf), "trailing bytes in the index file"
    h = np.concatenate(H) if H else np.zeros(0, np.uint64)
    c = np.concatenate(C) if C else np.zeros(0, np.int64)
    v = np.concatenate(V) if V else np.zeros(0, np.uint64)
    order = np.argsort(h, kind="stable")
    offs = np.concatenate(([0], np.cumsum(c)))
    idx = np.concatenate([np.arange(offs[j], offs[j + 1]) for j in order]) if len(order) else np.zeros(0, np.int64)
    keys = h[order]; cnt = c[order]
    return dict(w=w, e=e, n=n, q=q, lq=lq, k=k, flag=flag, n_seq=n_seq, names=names, seq_len=np.array(lens, np.uint32), fwd=fwd, rev=rev,
                keys=np.ascontiguousarray(keys), key_off=np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64), vals=np.ascontiguousarray(v[idx]))


class _OraIndex(C.Structure):
    _fields_ = [("w", C.c_int32), ("e", C.c_int32), ("n", C.c_int32), ("q", C.c_int32), ("lq", C.c_int32), ("k", C.c_int32),
                ("n_seq", C.c_uint32), ("seq_len", C.c_void_p), ("fwd", C.c_void_p), ("rev", C.c_void_p),
                ("n_keys", C.c_uint64), ("keys", C.c_void_p), ("key_off", C.c_void_p), ("vals", C.c_void_p)]


class MapOracle:
    """ctypes binding of the map-loop restatement (ra_oracle_map.c) over an index parsed by parse_index_file."""

    def __init__(self, index: dict):
        self.lib = L = C.CDLL(build_oracle())
        self.index = index
        self._fwd = (C.c_void_p * index["n_seq"])(*[a.ctypes.data for a in index["fwd"]])
        self._rev = (C.c_void_p * index["n_seq"])(*[a.ctypes.data for a in index["rev"]])
        self.ix = _OraIndex(index["w"], index["e"], index["n"], index["q"], index["lq"], index["k"], index["n_seq"],
                            index["seq_len"].ctypes.data, C.cast(self._fwd, C.c_void_p), C.cast(self._rev, C.c_void_p),
                            len(index["keys"]), index["keys"].ctypes.data, index["key_off"].ctypes.data, index["vals"].ctypes.data)
        vp = C.c_void_p
        L.ora_seq_to_sig.restype = C.c_uint32; L.ora_seq_to_sig.argtypes = [vp, C.c_int, _f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.ora_sketch.restype = C.c_uint32
        L.ora_sketch.argtypes = [_f32p, C.c_uint32] + [C.c_int] * 6 + [_u64p, _u32p, C.POINTER(C.c_int)]
        L.ora_reader_new.restype = vp; L.ora_reader_new.argtypes = [C.POINTER(_OraIndex), C.POINTER(RiMapOpt), C.c_int]
        L.ora_reader_free.argtypes = [vp]
        L.ora_reader_chunk.restype = C.c_int; L.ora_reader_chunk.argtypes = [vp, _f32p, C.c_uint32]
        L.ora_reader_chunk_events.restype = C.c_int; L.ora_reader_chunk_events.argtypes = [vp, _f32p, C.c_uint32]
        L.ora_reader_state.restype = C.c_uint32
        L.ora_reader_state.argtypes = [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.ora_reader_events.restype = C.POINTER(C.c_float); L.ora_reader_events.argtypes = [vp]
        L.ora_reader_chains.restype = C.c_uint32; L.ora_reader_chains.argtypes = [vp, C.c_int, vp, C.c_uint32, vp, C.c_uint64]
        L.ora_map_read.argtypes = [C.POINTER(_OraIndex), C.POINTER(RiMapOpt), C.c_int, _f32p, C.c_uint32, vp]
        L.ora_paf_format.restype = C.c_int; L.ora_paf_format.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_uint32, C.c_char_p, C.c_int]

    @staticmethod
    def mapopt(**kw) -> RiMapOpt:
        """ri_mapopt_init, roptions.c:5-61."""
        o = RiMapOpt(bp_per_sec=450, sample_rate=4000, chunk_size=4000, max_gap_length=2000, max_target_gap_length=5000,
                     chaining_band_length=5000, max_num_skips=25, min_num_anchors=2, num_best_chains=3, min_chaining_score=10.0,
                     step_size=1, min_events=50, max_num_chunk=30, min_chain_anchor=2, min_chain_anchor_out=2,
                     min_bestmap_ratio=1.2, min_bestmap_ratio_out=1.2, min_meanmap_ratio=5, min_meanmap_ratio_out=5,
                     mini_batch_size=500000000, window_length1=3, window_length2=6, threshold1=4.30265, threshold2=2.57058, peak_height=1.0,
                     t_threshold=1.5, tn_samples=5, ttest_freq=500, tmin_reads=500,
                     dtw_border_constraint=1, dtw_fill_method=1, dtw_band_radius_frac=0.10, dtw_match_bonus=0.4, dtw_min_score=20.0)
        for k, v in kw.items():
            setattr(o, k, v)
        return o

    def seq_to_sig(self, seq2bit, pore_vals, k=6, strand=0, contract=1):
        seq2bit = np.ascontiguousarray(seq2bit, np.uint8)
        out = np.zeros(len(seq2bit) + 1, np.float32)
        n = self.lib.ora_seq_to_sig(seq2bit.ctypes.data, len(seq2bit), _f32(pore_vals), k, strand, contract, out)
        return out[:n].copy()

    def sketch(self, ev):
        ev = _f32(ev); ix = self.index
        h = np.empty(len(ev) + 1, np.uint64); p = np.empty(len(ev) + 1, np.uint32); un = C.c_int(0)
        m = self.lib.ora_sketch(ev, len(ev), ix["w"], ix["e"], ix["n"], ix["q"], ix["lq"], ix["k"], h, p, C.byref(un))
        if un.value:
            raise NotImplementedError("minimizer / BLEND sketching is not restated")
        return h[:m].copy(), p[:m].copy()

    def idx_get(self, h: int) -> np.ndarray:
        k = np.searchsorted(self.index["keys"], np.uint64(h))
        if k < len(self.index["keys"]) and self.index["keys"][k] == np.uint64(h):
            return self.index["vals"][int(self.index["key_off"][k]):int(self.index["key_off"][k + 1])].copy()
        return np.zeros(0, np.uint64)

    def _chains(self, h, which):
        ch = np.zeros(256, CHAIN_OUT_DTYPE); an = np.zeros(1 << 17, ANCHOR_DTYPE)
        n = self.lib.ora_reader_chains(h, which, ch.ctypes.data, len(ch), an.ctypes.data, len(an))
        tot = int(ch["n_anchors"][:n].sum())
        return ch[:n].copy(), an[:tot].copy()

    def read_chains(self, opt: RiMapOpt, sig, max_chunks=None, stop_when_mapped=True, contract=1, events=None):
        """Chunk loop of map_worker_for for one read; per chunk: reg->chains, the pre-filter candidates in evaluation order, mapped.
        events = (ev, off): feed already detected per-chunk events instead of detecting from the signal."""
        sig = _f32(sig)
        h = self.lib.ora_reader_new(C.byref(self.ix), C.byref(opt), contract)
        out = []
        cs = int(opt.chunk_size); mc = int(max_chunks or opt.max_num_chunk)
        try:
            c = 0; start = 0
            while start < len(sig) and c < mc:
                end = min(len(sig), start + cs)
                if events is None:
                    mapped = self.lib.ora_reader_chunk(h, sig[start:end], end - start)
                else:
                    ev, off = events
                    e = _f32(ev[int(off[c]):int(off[c + 1])])
                    mapped = self.lib.ora_reader_chunk_events(h, e if len(e) else np.zeros(1, np.float32), len(e))
                ch, an = self._chains(h, 0); cand, can = self._chains(h, 1)
                ne = C.c_uint32(0); ro = C.c_uint32(0); cells = C.c_uint64(0); ns = C.c_uint64(0); nh = C.c_uint64(0)
                self.lib.ora_reader_state(h, C.byref(ne), C.byref(ro), C.byref(cells), C.byref(ns), C.byref(nh))
                out.append(dict(chains=ch, anchors=an, cand=cand, cand_anchors=can, mapped=bool(mapped), n_events=ne.value, offset=ro.value,
                                cells=cells.value, seeds=ns.value, hits=nh.value))
                if mapped and stop_when_mapped:
                    break
                start += cs; c += 1
            n = out[-1]["n_events"] if out else 0
            ev = np.ctypeslib.as_array(self.lib.ora_reader_events(h), (max(n, 1),))[:n].copy() if n else np.zeros(0, np.float32)
            return out, ev
        finally:
            self.lib.ora_reader_free(h)

    def map_reads(self, opt: RiMapOpt, sig, sig_off, l_sig, contract=1, names=None):
        sig = _f32(sig)
        res = np.zeros(len(l_sig), PAF_DTYPE)
        for r in range(len(l_sig)):
            a = int(sig_off[r]); n = int(l_sig[r])
            self.lib.ora_map_read(C.byref(self.ix), C.byref(opt), contract, sig[a:a + n] if n else np.zeros(1, np.float32), n, res[r:r + 1].ctypes.data)
        return res

    def paf_lines(self, res, names=None):
        out = []
        buf = C.create_string_buffer(4096)
        for r in range(len(res)):
            nm = (names[r] if names else f"read_{r}").encode()
            rid = int(res["ref_id"][r])
            n = self.lib.ora_paf_format(res[r:r + 1].ctypes.data, nm, self.index["names"][rid].encode(), int(self.index["seq_len"][rid]), buf, 4096)
            out.append(buf.raw[:n].decode())
        return out


def build_flat_index(seqs, pore_vals, preset="sensitive", k=6, contract=1, names=None) -> dict:
    """Index construction restated (rawindex.cpp:127-153 per sequence: ri_seq_to_sig + ri_sketch for strand flag 1 then 0;
    ri_idx_add / worker_post: group by hash, positions ascending): the same flat table parse_index_file returns for an index
    the reference built.  Signal arrays are seq.len floats with a zero tail, as the reference allocates and dumps them
    (rawindex.cpp:136-137, 294-295)."""
    e, q, lq, w, n = PRESETS[preset]
    mo = MapOracle.__new__(MapOracle)
    mo.lib = C.CDLL(build_oracle())
    mo.lib.ora_seq_to_sig.restype = C.c_uint32
    mo.lib.ora_seq_to_sig.argtypes = [C.c_void_p, C.c_int, _f32p, C.c_int, C.c_int, C.c_int, _f32p]
    mo.lib.ora_sketch.restype = C.c_uint32
    mo.lib.ora_sketch.argtypes = [_f32p, C.c_uint32] + [C.c_int] * 6 + [_u64p, _u32p, C.POINTER(C.c_int)]
    mo.index = dict(w=w, e=e, n=n, q=q, lq=lq, k=k)
    H, Y, fwd, rev = [], [], [], []
    for i, s in enumerate(seqs):
        for strand, store in ((1, fwd), (0, rev)):
            sig = mo.seq_to_sig(s, pore_vals, k, strand, contract)
            full = np.zeros(len(s), np.float32); full[:len(sig)] = sig
            store.append(full)
            h, p = mo.sketch(sig)
            H.append(h); Y.append((np.uint64(i) << np.uint64(32)) | (p.astype(np.uint64) << np.uint64(1)) | np.uint64(strand))
    h = np.concatenate(H); y = np.concatenate(Y)
    order = np.lexsort((y, h))
    h = h[order]; y = y[order]
    keys, first = np.unique(h, return_index=True)
    key_off = np.concatenate((first, [len(h)])).astype(np.uint64)
    return dict(w=w, e=e, n=n, q=q, lq=lq, k=k, flag=0, n_seq=len(seqs), names=names or [f"ref{i}" for i in range(len(seqs))],
                seq_len=np.array([len(s) for s in seqs], np.uint32), fwd=fwd, rev=rev, keys=np.ascontiguousarray(keys), key_off=key_off,
                vals=np.ascontiguousarray(y))


def write_index_file(path: str, ix: dict, names=None) -> None:
    """Writes a flat table in the reference's index file format (ri_idx_dump, rawindex.cpp:275-315), so that index loaders
    (the reference's ri_idx_load, the product's ra_index_load_file) can be driven where the reference build is absent."""
    b = 14
    keys = ix["keys"]; off = ix["key_off"].astype(np.int64); vals = ix["vals"]
    bucket = (keys & np.uint64((1 << b) - 1)).astype(np.int64)
    order = np.argsort(bucket, kind="stable")
    bstart = np.searchsorted(bucket[order], np.arange((1 << b) + 1))
    with open(path, "wb") as f:
        f.write(b"RI")
        f.write(np.array([ix["w"], ix["e"], ix["n"], ix["q"], ix["lq"], ix["k"], ix["n_seq"], ix.get("flag", 0)], np.uint32).tobytes())
        nm = names or ix.get("names") or [f"ref{i}" for i in range(ix["n_seq"])]
        for i in range(ix["n_seq"]):
            f.write(bytes([len(nm[i])])); f.write(nm[i].encode())
            f.write(np.uint32(ix["seq_len"][i]).tobytes())
            f.write(np.ascontiguousarray(ix["fwd"][i], np.float32).tobytes()); f.write(np.ascontiguousarray(ix["rev"][i], np.float32).tobytes())
        for i in range(1 << b):
            sel = order[bstart[i]:bstart[i + 1]]
            cnt = (off[sel + 1] - off[sel])
            multi = cnt > 1
            p = np.concatenate([vals[off[j]:off[j + 1]] for j in sel[multi]]) if multi.any() else np.zeros(0, np.uint64)
            f.write(np.int32(len(p)).tobytes()); f.write(p.tobytes())
            f.write(np.uint32(len(sel)).tobytes())
            if len(sel) == 0:
                continue
            kv = np.zeros((len(sel), 2), np.uint64)
            kv[:, 0] = ((keys[sel] >> np.uint64(b)) << np.uint64(1)) | (~multi).astype(np.uint64)
            pstart = np.concatenate(([0], np.cumsum(cnt[multi])))[:-1]
            v = np.zeros(len(sel), np.uint64)
            v[~multi] = vals[off[sel[~multi]]]
            v[multi] = (pstart.astype(np.uint64) << np.uint64(32)) | cnt[multi].astype(np.uint64)
            kv[:, 1] = v
            f.write(kv.tobytes())
