"""CPU: the C-ABI library loads and exports every symbol include/rawalign_b200.h declares; host-only
entry points behave; with no GPU the product fails loudly instead of falling back."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "rawalign_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ra_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from rawalign_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/rawalign_b200.h but not exported"
    assert set(syms) == set(_lib.EXPORTED_SYMBOLS)


def test_mapopt_defaults_match_roptions():
    import rawalign_b200 as ra
    o = ra.ri_mapopt_init()
    assert (o.window_length1, o.window_length2) == (3, 6)
    assert np.float32(o.threshold1) == np.float32(4.30265) and np.float32(o.threshold2) == np.float32(2.57058)
    assert o.peak_height == 1.0 and o.dtw_border_constraint == 1 and o.dtw_fill_method == 1
    assert np.float32(o.dtw_band_radius_frac) == np.float32(0.10) and np.float32(o.dtw_match_bonus) == np.float32(0.4)
    assert o.dtw_min_score == 20.0 and o.fp_contract == 1
    assert C.sizeof(o) == 48


def test_cell_count_matches_oracle(oracle):
    import rawalign_b200 as ra
    rng = np.random.default_rng(3)
    for _ in range(500):
        al, bl, R = int(rng.integers(1, 300)), int(rng.integers(1, 300)), int(rng.integers(0, 20))
        assert ra.Context.dtw_cells(al, bl, R) == oracle.dtw_cells(al, bl, R)
    assert ra.Context.dtw_cells(30, 40, -1) == 1200


def test_struct_layouts():
    import rawalign_b200 as ra
    assert ra.ITEM_DTYPE.itemsize == 32 and ra.ANCHOR_DTYPE.itemsize == 8 and ra.CHAIN_DTYPE.itemsize == 32


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import rawalign_b200 as ra
    with pytest.raises(ra.RawAlignError) as e:
        ra.Context(0)
    assert e.value.code == -1          # RA_ENODEV
