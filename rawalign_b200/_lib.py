"""ctypes binding of librawalign_b200.so (the C ABI declared in include/rawalign_b200.h).

There is no fallback: if the shared library is missing or no sm_100 GPU is present, importing the
binding or creating a context raises.  Build the library with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C rawalign_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librawalign_b200.so")

RA_OK, RA_ENODEV, RA_ECUDA, RA_EINVAL, RA_ENOMEM, RA_EUNSUP = 0, -1, -2, -3, -4, -5

ITEM_DTYPE = np.dtype([("a_off", "<u8"), ("b_off", "<u8"), ("a_len", "<u4"), ("b_len", "<u4"),
                       ("band_radius", "<i4"), ("exclude_last", "<u4")])
ANCHOR_DTYPE = np.dtype([("target_position", "<u4"), ("query_position", "<u4")])
CHAIN_DTYPE = np.dtype([("anchor_off", "<u8"), ("n_anchors", "<u4"), ("group", "<u4"),
                        ("read_off", "<u8"), ("ref_off", "<u8")])


class MapOpt(C.Structure):
    """ra_mapopt_t: the hot-path fields of ri_mapopt_t (reference roptions.h:39-87)."""
    _fields_ = [("window_length1", C.c_uint32), ("window_length2", C.c_uint32),
                ("threshold1", C.c_float), ("threshold2", C.c_float), ("peak_height", C.c_float),
                ("dtw_border_constraint", C.c_uint32), ("dtw_fill_method", C.c_uint32),
                ("dtw_band_radius_frac", C.c_float), ("dtw_match_bonus", C.c_float), ("dtw_min_score", C.c_float),
                ("fp_contract", C.c_uint32), ("reserved", C.c_uint32)]


class RiMapOpt(C.Structure):
    """ra_ri_mapopt_t = ri_mapopt_t (reference roptions.h:39-87), field for field."""
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


class MapStats(C.Structure):
    """ra_map_stats_t"""
    _fields_ = [(n, C.c_float) for n in ("decode_ms", "events_ms", "seed_ms", "sort_ms", "chain_ms", "dtw_ms", "select_ms", "total_ms")] + \
               [("rounds", C.c_uint32), ("reserved", C.c_uint32)] + \
               [(n, C.c_uint64) for n in ("chunks", "events", "seeds", "anchors", "chains_evaluated", "launches", "cells_issued", "cells_useful", "h2d_bytes", "d2h_bytes")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


PAF_DTYPE = np.dtype([("mapped", "<u4"), ("ref_id", "<u4"), ("read_length", "<u4"), ("read_start_position", "<u4"),
                      ("read_end_position", "<u4"), ("fragment_start_position", "<u4"), ("fragment_length", "<u4"), ("mapq", "<u4"),
                      ("rev", "<u4"), ("ci", "<u4"), ("sl", "<u4"), ("cm", "<u4"), ("nc", "<u4"),
                      ("s1", "<f4"), ("s2", "<f4"), ("sm", "<f4"), ("at", "<f4"), ("aq", "<f4"), ("alns", "<f4"),
                      ("n_events", "<u4"), ("reserved", "<u4")])


class RawAlignError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"rawalign_b200 error {code}: {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it (make -C rawalign_b200/csrc); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    L.ra_last_error.restype = C.c_char_p
    L.ra_init.argtypes = [i32, C.POINTER(vp)]
    L.ra_destroy.argtypes = [vp]; L.ra_destroy.restype = None
    L.ra_mapopt_init.argtypes = [C.POINTER(MapOpt)]; L.ra_mapopt_init.restype = None
    L.ra_set_opt.argtypes = [vp, C.POINTER(MapOpt)]
    L.ra_launch_count.argtypes = [vp]; L.ra_launch_count.restype = u64
    L.ra_last_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.ra_ref_load.argtypes = [vp, u32, vp, vp, vp, vp]
    L.ra_ref_load_flat.argtypes = [vp, vp, u64]
    L.ra_ref_device.argtypes = [vp, C.POINTER(vp), C.POINTER(u64)]
    L.ra_ref_alloc.argtypes = [vp, u64]
    L.ra_detect_events_batch.argtypes = [vp, vp, u64, vp, vp, u32, vp, u64, vp]
    L.ra_detect_events_batch_dev.argtypes = [vp, vp, vp, vp, u32, u32, vp, u64, vp, C.POINTER(u64)]
    L.ra_detect_events_raw_batch.argtypes = [vp, vp, u64, vp, vp, vp, vp, u32, u32, u32, vp, vp, vp, u64, vp]
    L.ra_last_event_positions.argtypes = [vp, vp, u64]
    L.ra_dtw_batch.argtypes = [vp, vp, u64, vp, u64, vp, u32, vp]
    L.ra_dtw_batch_dev.argtypes = [vp, vp, vp, vp, vp, u32, vp]
    L.ra_dtw_cells.argtypes = [u32, u32, C.c_int32]; L.ra_dtw_cells.restype = u64
    L.ra_dtw_tb_batch.argtypes = [vp, vp, u64, vp, u64, vp, u32, vp, vp, vp, vp, u64, vp]
    L.ra_align_chains_batch.argtypes = [vp, vp, u64, vp, u64, vp, u64, vp, u32, i32, vp, vp, vp, C.POINTER(u64), C.POINTER(u64)]
    L.ra_align_chains_batch_dev.argtypes = [vp, vp, vp, vp, u64, vp, u32, vp, u32, i32, vp, vp, vp, C.POINTER(u64), C.POINTER(u64)]
    L.ra_align_chains_cigar_batch.argtypes = [vp, vp, u64, vp, u64, vp, u64, vp, u32, vp, vp, vp, vp, vp, u64, vp]
    L.ra_get_opt.argtypes = [vp, C.POINTER(MapOpt)]
    L.ra_set_tuning.argtypes = [vp, C.c_char_p, C.c_int64]; L.ra_set_tuning.restype = C.c_int
    L.ra_ri_mapopt_init.argtypes = [C.POINTER(RiMapOpt)]; L.ra_ri_mapopt_init.restype = None
    L.ra_index_load_file.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.ra_index_load_flat.argtypes = [vp] + [C.c_int32] * 6 + [u32, vp, vp, vp, u64, vp, vp, vp, C.POINTER(vp)]
    L.ra_index_destroy.argtypes = [vp]; L.ra_index_destroy.restype = None
    L.ra_index_info.argtypes = [vp, C.POINTER(C.c_int32), vp]
    L.ra_map_batch.argtypes = [vp, vp, C.POINTER(RiMapOpt), vp, vp, vp, vp, vp, vp, vp, vp, u32, vp, C.POINTER(MapStats)]
    L.ra_map_last_cigar.argtypes = [vp, u32, vp, vp, vp, vp, u64]
    L.ra_index_tables.argtypes = [vp, vp, vp]; L.ra_index_tables.restype = C.c_int
    L.ra_index_alloc.argtypes = [vp, C.POINTER(C.c_int32), u64, u64, vp, C.POINTER(vp)]; L.ra_index_alloc.restype = C.c_int
    L.ra_index_clone.argtypes = [vp, vp, C.POINTER(vp)]; L.ra_index_clone.restype = C.c_int
    L.ra_detect_events_noalloc.argtypes = [vp, vp, vp, vp, u32, u32, vp, u64, vp, vp]; L.ra_detect_events_noalloc.restype = C.c_int
    L.ra_index_build.argtypes = [vp] + [C.c_int32] * 4 + [u32, vp, vp, vp, C.POINTER(vp)]; L.ra_index_build.restype = C.c_int
    L.ra_index_export.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), vp, vp, vp]; L.ra_index_export.restype = C.c_int
    for name in ("ra_get_opt", "ra_index_load_file", "ra_index_load_flat", "ra_index_info", "ra_map_batch", "ra_map_last_cigar"):
        getattr(L, name).restype = C.c_int
    for name in ("ra_init", "ra_set_opt", "ra_last_timing", "ra_ref_load", "ra_ref_load_flat", "ra_ref_device", "ra_ref_alloc",
                 "ra_detect_events_batch", "ra_detect_events_batch_dev", "ra_detect_events_raw_batch", "ra_last_event_positions", "ra_dtw_batch", "ra_dtw_batch_dev",
                 "ra_dtw_tb_batch", "ra_align_chains_batch", "ra_align_chains_batch_dev", "ra_align_chains_cigar_batch"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != RA_OK:
        raise RawAlignError(rc, load().ra_last_error().decode(errors="replace"))


EXPORTED_SYMBOLS = [
    "ra_init", "ra_destroy", "ra_last_error", "ra_mapopt_init", "ra_set_opt", "ra_launch_count", "ra_last_timing",
    "ra_ref_load", "ra_ref_load_flat", "ra_ref_device", "ra_ref_alloc",
    "ra_detect_events_batch", "ra_detect_events_batch_dev", "ra_detect_events_raw_batch", "ra_last_event_positions",
    "ra_dtw_batch", "ra_dtw_batch_dev", "ra_dtw_cells", "ra_dtw_tb_batch",
    "ra_align_chains_batch", "ra_align_chains_batch_dev", "ra_align_chains_cigar_batch",
    "ra_get_opt", "ra_set_tuning", "ra_ri_mapopt_init", "ra_index_load_file", "ra_index_load_flat", "ra_index_destroy", "ra_index_info", "ra_map_batch", "ra_map_last_cigar", "ra_index_tables", "ra_index_alloc",
    "ra_index_build", "ra_index_export", "ra_index_clone", "ra_detect_events_noalloc",
]
