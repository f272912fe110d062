"""CPU: randomized differential test of the oracle against the reference's own objects
(oracle/_ref, built where /root/reference exists).  Skipped where the reference build is absent."""
import numpy as np
import pytest

from conftest import bits
from oracle import Reference

pytestmark = pytest.mark.skipif(not Reference.available(True), reason="oracle/_ref not built (no /root/reference here)")


def test_dtw_differential(oracle):
    ref = Reference(True)
    rng = np.random.default_rng(11)
    for it in range(4000):
        al, bl, R = int(rng.integers(1, 150)), int(rng.integers(1, 150)), int(rng.integers(0, 16))
        a = rng.uniform(-2.5, 2.5, al).astype(np.float32); b = rng.uniform(-2.5, 2.5, bl).astype(np.float32)
        ex = bool(rng.integers(0, 2))
        assert bits(oracle.dtw_banded(a, b, R, ex)) == bits(ref.dtw_banded(a, b, R, ex)), (al, bl, R, ex)
        if it % 8 == 0:
            assert bits(oracle.dtw_global(a, b, ex)) == bits(ref.dtw_global(a, b, ex))
            c1, i1, j1, d1 = oracle.dtw_global_tb(a, b, ex); c2, i2, j2, d2 = ref.dtw_global_tb(a, b, ex)
            assert bits(c1) == bits(c2) and np.array_equal(i1, i2) and np.array_equal(j1, j2) and np.array_equal(bits(d1), bits(d2))


def test_events_differential(oracle):
    from oracle.gen_golden import sim_signal
    rng = np.random.default_rng(12)
    for it in range(150):
        n = 4000 if it % 3 else int(rng.integers(1, 4000))
        s = sim_signal(n + 40, rng)[:n]
        for contract, fma in ((1, True), (0, False)):
            got = oracle.detect_events(s, contract=contract); want = Reference(fma).detect_events(s)
            assert len(got) == len(want) and np.array_equal(bits(got), bits(want)), (it, n, contract)


def test_check_dtw_generator_matches_std_random():
    """ref_checkdtw_vector reproduces check_dtw.cpp:13-27 (mt19937 + uniform_real_distribution<double>);
    numpy's MT19937 is the same generator: two 32-bit draws per double, (lo + hi*2^32) / 2^64."""
    ref = Reference(True)
    for seed in (0, 1, 1428, 99999):
        bg = np.random.MT19937(); bg._legacy_seeding(seed)
        raw = np.random.Generator(bg).bit_generator.random_raw(64).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        lo, hi = raw[0::2].astype(np.float64), raw[1::2].astype(np.float64)
        u = (lo + hi * 4294967296.0) / 18446744073709551616.0
        want = (u * 5.0 - 2.5).astype(np.float32)
        got = ref.checkdtw_vector(32, seed)
        assert np.array_equal(got, want)
