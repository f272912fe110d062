"""rawalign_b200 -- B200 (sm_100a) implementation of RawAlign's data-parallel hot path:
batched event detection (reference src/revent.c) and batched DTW evaluation of chained candidate
regions (reference src/dtw.cpp, src/rmap.cpp:align_chain), behind the C ABI in
include/rawalign_b200.h.  ``csrc/`` holds the CUDA kernels and the C ABI, ``api.py`` mirrors the
reference's function names on top of it, the seeded synthetic workload generator lives outside the package (``workload/synth.py``).
"""
from ._lib import ANCHOR_DTYPE, CHAIN_DTYPE, ITEM_DTYPE, LIB_PATH, MapOpt, RawAlignError  # noqa: F401
from .api import (  # noqa: F401
    RI_M_DTW_BORDER_CONSTRAINT_GLOBAL, RI_M_DTW_BORDER_CONSTRAINT_LOCAL, RI_M_DTW_BORDER_CONSTRAINT_SPARSE,
    RI_M_DTW_FILL_METHOD_BANDED, RI_M_DTW_FILL_METHOD_FULL, Context, make_items, ri_mapopt_init,
)
from ._lib import PAF_DTYPE, MapStats, RiMapOpt  # noqa: F401,E402
from .mapper import RI_M_DTW_EVALUATE_CHAINS, RI_M_DTW_OUTPUT_CIGAR, Index, ri_mapopt_full  # noqa: F401,E402
