"""CPU: the oracle (oracle/ra_oracle.c) against the committed fixtures that were generated from the
reference's own dtw.cpp / revent.c objects (oracle/gen_golden.py).  Bit-exact."""
import numpy as np

from conftest import bits, unpack


def test_banded_cost_bit_exact(oracle, dtw_golden):
    g = dtw_golden
    n = len(g["band_radius"])
    bad = 0
    for k in range(n):
        a = unpack(g["a"], g["a_off"], k); b = unpack(g["b"], g["b_off"], k)
        got = oracle.dtw_banded(a, b, int(g["band_radius"][k]), bool(g["exclude_last"][k]))
        bad += bits(got) != bits(g["banded_cost"][k])
    assert bad == 0


def test_full_cost_bit_exact(oracle, dtw_golden):
    g = dtw_golden
    for k in range(0, len(g["band_radius"]), 2):
        a = unpack(g["a"], g["a_off"], k); b = unpack(g["b"], g["b_off"], k)
        got = oracle.dtw_global(a, b, bool(g["exclude_last"][k]))
        assert bits(got) == bits(g["full_cost"][k]), k


def test_traceback_identical(oracle, dtw_golden):
    g = dtw_golden
    for t, k in enumerate(g["tb_idx"]):
        a = unpack(g["a"], g["a_off"], k); b = unpack(g["b"], g["b_off"], k)
        cost, pi, pj, pd = oracle.dtw_global_tb(a, b, bool(g["exclude_last"][k]))
        assert bits(cost) == bits(g["tb_cost"][t])
        assert np.array_equal(pi, unpack(g["tb_i"], g["tb_off"], t))
        assert np.array_equal(pj, unpack(g["tb_j"], g["tb_off"], t))
        assert np.array_equal(bits(pd), bits(unpack(g["tb_diff"], g["tb_off"], t)))


def test_events_bit_exact_both_contraction_modes(oracle, event_golden):
    g = event_golden
    n = len(g["sig_off"]) - 1
    for k in range(n):
        s = unpack(g["sig"], g["sig_off"], k)
        for contract, key in ((1, "ev_fma"), (0, "ev_nofma")):
            want = unpack(g[key], g[key + "_off"], k)
            got = oracle.detect_events(s, contract=contract)
            assert len(got) == len(want), (k, contract)
            assert np.array_equal(bits(got), bits(want)), (k, contract)


def test_cell_count_closed_form(oracle):
    rng = np.random.default_rng(5)
    for _ in range(300):
        al, bl, R = int(rng.integers(1, 200)), int(rng.integers(1, 200)), int(rng.integers(0, 12))
        a = rng.normal(size=al).astype(np.float32); b = rng.normal(size=bl).astype(np.float32)
        _, cells = oracle.dtw_banded(a, b, R, return_cells=True)
        assert cells == oracle.dtw_cells(al, bl, R)
        assert cells <= al * bl
    assert oracle.dtw_cells(7, 9, -1) == 63


def test_wide_band_equals_full(oracle):
    """A band that covers the whole matrix must reproduce DTW_global bit for bit (the property
    check_dtw.cpp:128-163 relies on)."""
    rng = np.random.default_rng(6)
    for _ in range(100):
        al, bl = int(rng.integers(1, 60)), int(rng.integers(1, 60))
        a = rng.normal(size=al).astype(np.float32); b = rng.normal(size=bl).astype(np.float32)
        assert bits(oracle.dtw_banded(a, b, 2 * max(al, bl))) == bits(oracle.dtw_global(a, b))


def _chain(rng, n_anchors, q0=5, t0=11):
    from oracle import ANCHOR_DTYPE
    dq = 1 + rng.geometric(0.25, n_anchors - 1); dt = np.maximum(1, dq + rng.integers(-1, 2, n_anchors - 1))
    q = q0 + np.concatenate(([0], np.cumsum(dq))); t = t0 + np.concatenate(([0], np.cumsum(dt)))
    a = np.zeros(n_anchors, ANCHOR_DTYPE)
    a["query_position"] = q[::-1]; a["target_position"] = t[::-1]
    return a


def test_align_chain_restates_rmap(oracle):
    """align_chain (rmap.cpp:181-313): sparse = ordered sum of anchor-to-anchor costs with
    exclude_last on all but the final part; global = one problem; early termination -> -1e10."""
    from oracle import DtwOpt
    rng = np.random.default_rng(8)
    for _ in range(50):
        anc = _chain(rng, int(rng.integers(2, 30)))
        read = rng.normal(size=int(anc["query_position"][0]) + 3).astype(np.float32)
        ref = rng.normal(size=int(anc["target_position"][0]) + 3).astype(np.float32)
        for fill in (0, 1):
            opt = DtwOpt(1, fill, 0.1, 0.4, 20.0)
            cost = np.float32(0); n_al = 0
            parts = len(anc) - 1
            for p in range(parts):
                s, e = anc[parts - p], anc[parts - p - 1]
                ra = read[s["query_position"]:e["query_position"] + 1]; rb = ref[s["target_position"]:e["target_position"] + 1]
                excl = p != parts - 1
                sub = oracle.dtw_global(ra, rb, excl) if fill == 0 else oracle.dtw_banded(ra, rb, max(1, int(np.float32(len(ra)) * np.float32(0.1))), excl)
                cost = np.float32(cost + sub); n_al += len(ra)
            want = np.float32(np.float32(np.float32(n_al) * np.float32(0.4)) - cost)
            assert bits(oracle.align_chain(anc, read, ref, opt)) == bits(want)
            # a min_score nothing can reach drops the chain before any work
            s, cells = oracle.align_chain(anc, read, ref, opt, min_score=1e9, return_cells=True)
            assert s == np.float32(-1e10) and cells == 0
            og = DtwOpt(0, fill, 0.1, 0.4, 20.0)
            s0, e0 = anc[-1], anc[0]
            ra = read[s0["query_position"]:e0["query_position"] + 1]; rb = ref[s0["target_position"]:e0["target_position"] + 1]
            c = oracle.dtw_global(ra, rb) if fill == 0 else oracle.dtw_banded(ra, rb, max(1, int(np.float32(len(ra)) * np.float32(0.1))))
            assert bits(oracle.align_chain(anc, read, ref, og)) == bits(np.float32(np.float32(np.float32(len(ra)) * np.float32(0.4)) - c))


def test_decode_raw_restates_rsig(oracle):
    """ri_read_sig's conversion (rsig.cpp:213-222) against an independent numpy fp32 evaluation, filter included."""
    rng = np.random.default_rng(5)
    raw = rng.integers(-200, 2500, 20000).astype(np.int16)
    for offset, scale in ((10.0, 1467.61 / 8192.0), (-3.5, 0.21), (0.0, 0.05)):
        o32, s32 = np.float32(offset), np.float32(scale)
        pa = (raw.astype(np.float32) + o32) * s32
        want = pa[(pa > 30) & (pa < 200)]
        got = oracle.decode_raw(raw, o32, s32)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert len(oracle.decode_raw(np.zeros(0, np.int16), 1.0, 1.0)) == 0
