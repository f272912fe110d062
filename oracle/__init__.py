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
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
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
