"""CPU: the 3-instruction division by the detector window used by the fused event kernel
(rawalign_b200/csrc/events_fused.cu:fdiv_w) is the correctly rounded quotient the reference computes
(revent.c:59-61, :69) -- exhaustive over all 2^23 mantissas, three exponents, both signs, for every
window the fused kernel accepts."""
import ctypes as C

import pytest


@pytest.mark.parametrize("w", [2, 3, 5, 6, 7, 9, 11, 13, 14])
def test_fast_division_by_window_is_exact(w):
    from oracle import build_oracle
    L = C.CDLL(build_oracle())
    L.ora_check_fastdiv.restype = C.c_uint64
    L.ora_check_fastdiv.argtypes = [C.c_uint32]
    assert L.ora_check_fastdiv(w) == 0
