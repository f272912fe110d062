"""GPU (-m gpu): the CUDA path through the C ABI against the oracle and the committed reference
fixtures.  Everything is asserted BIT-EXACT (integer compare of the float bit patterns)."""
import numpy as np
import pytest

from conftest import bits, unpack

pytestmark = pytest.mark.gpu


def _items(ra, a_lens, b_lens, R, ex):
    a_off = np.concatenate(([0], np.cumsum(a_lens)[:-1])); b_off = np.concatenate(([0], np.cumsum(b_lens)[:-1]))
    return ra.make_items(a_off, b_off, a_lens, b_lens, R, ex)


def test_extension_is_loaded(ctx):
    import os
    import rawalign_b200 as ra
    maps = open("/proc/self/maps").read()
    assert os.path.basename(ra.LIB_PATH) in maps, "native library not loaded"


# ---- K1 ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("unfused", [0, 1, 2])   # K1 variant: 0 = v4 pipeline (default), 1 = v2 five kernels, 2 = v3 fused
@pytest.mark.parametrize("contract,key", [(1, "ev_fma"), (0, "ev_nofma")])
def test_events_golden(ctx, event_golden, contract, key, unfused):
    import rawalign_b200 as ra
    g = event_golden
    opt = ra.ri_mapopt_init(); opt.fp_contract = contract; opt.reserved = unfused
    ctx.set_opt(opt)
    try:
        n = len(g["sig_off"]) - 1
        # pad every chunk start to 16 bytes and also run an unaligned layout
        for pad in (4, 1):
            lens = np.diff(g["sig_off"]).astype(np.uint32)
            starts = np.zeros(n, np.uint64); pos = 3 if pad == 1 else 0
            for k in range(n):
                starts[k] = pos; pos += int(lens[k]); pos = (pos + pad - 1) // pad * pad
            sig = np.zeros(pos + 8, np.float32)
            for k in range(n):
                sig[int(starts[k]):int(starts[k]) + int(lens[k])] = unpack(g["sig"], g["sig_off"], k)
            ev, off = ctx.detect_events_batch(sig, starts, lens)
            assert np.array_equal(off.astype(np.int64), g[key + "_off"]), "event counts differ"
            assert np.array_equal(bits(ev), bits(g[key])), "event values differ"
    finally:
        ctx.set_opt(ra.ri_mapopt_init())


def test_events_vs_oracle_simulated_reads(ctx, oracle):
    from workload import synth
    ref = synth.random_reference(200_000, seed=1)
    reads = synth.simulate_reads(ref, 96, seed=21, max_chunks=8)
    coff, clen, _ = synth.chunk_table(reads.read_off, reads.read_len, max_chunks=8)
    ev, off = ctx.detect_events_batch(reads.sig, coff, clen)
    ev_o, off_o = oracle.detect_events_batch(reads.sig, coff, clen)
    assert np.array_equal(off, off_o)
    assert np.array_equal(bits(ev), bits(ev_o))
    per_chunk = np.diff(off.astype(np.int64))
    assert per_chunk[clen == 4000].mean() > 300       # pore-model-like signal: ~400-600 events per 4000 samples


def test_events_single_call_and_empty(ctx, oracle):
    rng = np.random.default_rng(4)
    s = np.clip(rng.normal(90, 13, 4000), 31, 199).astype(np.float32)
    assert np.array_equal(bits(ctx.detect_events(s)), bits(oracle.detect_events(s)))
    assert len(ctx.detect_events(np.full(7, 80.0, np.float32))) == 0
    ev, off = ctx.detect_events_batch(np.zeros(4, np.float32), np.zeros(0, np.uint64), np.zeros(0, np.uint32))
    assert len(ev) == 0 and off[0] == 0


@pytest.mark.parametrize("w1,w2,t1,t2", [(7, 14, 2.5, 9.0), (4, 9, 3.0, 2.0), (2, 5, 4.0, 2.5), (1, 6, 4.3, 2.6)])
def test_events_other_windows(ctx, oracle, w1, w2, t1, t2):
    """Detector windows are run-time options (RNA uses 7/14, roptions.c:56-60); w < 2 silences a track (revent.c:46)."""
    import rawalign_b200 as ra
    from oracle import EvOpt
    from oracle.gen_golden import sim_signal
    rng = np.random.default_rng(50 + w1)
    lens = np.array([4000, 4000, 1234, 27, 13, 4000], np.uint32)
    offs = np.zeros(len(lens), np.uint64); offs[1:] = np.cumsum((lens[:-1] + 3) // 4 * 4)
    sig = np.zeros(int(offs[-1]) + 4000, np.float32)
    for k in range(len(lens)):
        sig[int(offs[k]):int(offs[k]) + lens[k]] = sim_signal(int(lens[k]) + 40, rng)[:lens[k]]
    opt = ra.ri_mapopt_init(); opt.window_length1 = w1; opt.window_length2 = w2; opt.threshold1 = t1; opt.threshold2 = t2
    ctx.set_opt(opt)
    try:
        ev, off = ctx.detect_events_batch(sig, offs, lens)
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    ev_o, off_o = oracle.detect_events_batch(sig, offs, lens, EvOpt(w1, w2, t1, t2, 1.0))
    assert np.array_equal(off, off_o) and np.array_equal(bits(ev), bits(ev_o))


def test_events_many_chunks_sub_batched(ctx, oracle):
    """More chunks than one launch group (32768): the CSR scan must chain across sub-batches."""
    rng = np.random.default_rng(77)
    n = 40_000
    lens = rng.integers(40, 120, n).astype(np.uint32)
    offs = np.zeros(n, np.uint64); offs[1:] = np.cumsum((lens[:-1] + 3) // 4 * 4)
    lv = np.repeat(rng.normal(90, 13, (int(offs[-1]) + 200) // 8 + 2), 8)[: int(offs[-1]) + 128]
    sig = np.clip(lv + rng.normal(0, 2, len(lv)), 31, 199).astype(np.float32)
    ev, off = ctx.detect_events_batch(sig, offs, lens)
    ev_o, off_o = oracle.detect_events_batch(sig, offs, lens)
    assert np.array_equal(off, off_o) and np.array_equal(bits(ev), bits(ev_o))
    assert off[-1] > n


# ---- K2 / K3 ------------------------------------------------------------------------------------------
def test_dtw_golden_banded_and_full(ctx, dtw_golden):
    import rawalign_b200 as ra
    g = dtw_golden
    n = len(g["band_radius"])
    a_len = np.diff(g["a_off"]); b_len = np.diff(g["b_off"])
    it = ra.make_items(g["a_off"][:-1], g["b_off"][:-1], a_len, b_len, g["band_radius"], g["exclude_last"])
    cost = ctx.dtw_batch(g["a"], g["b"], it)
    assert np.array_equal(bits(cost), bits(g["banded_cost"]))
    it["band_radius"] = -1
    cost = ctx.dtw_batch(g["a"], g["b"], it)
    assert np.array_equal(bits(cost), bits(g["full_cost"]))
    assert n > 1000


@pytest.mark.parametrize("generic", [0, 1])     # 0: register-wavefront kernel where the band fits a warp; 1: generic kernel for everything
def test_dtw_vs_oracle_random_and_wide(ctx, oracle, generic):
    import rawalign_b200 as ra
    opt = ra.ri_mapopt_init(); opt.reserved = 4 * generic
    ctx.set_opt(opt)
    rng = np.random.default_rng(31)
    shapes = [(int(rng.integers(1, 700)), int(rng.integers(1, 700)), int(rng.integers(0, 80))) for _ in range(300)]
    shapes += [(5000, 5200, 500), (12000, 11000, 1200), (2, 9000, 1), (9000, 3, 900), (3000, 3000, -1), (1, 1, 1), (1, 1, -1)]
    a_lens = np.array([s[0] for s in shapes]); b_lens = np.array([s[1] for s in shapes])
    R = np.array([s[2] for s in shapes]); ex = rng.integers(0, 2, len(shapes))
    A = rng.normal(0, 1, int(a_lens.sum())).astype(np.float32); B = rng.normal(0, 1, int(b_lens.sum())).astype(np.float32)
    it = _items(ra, a_lens, b_lens, R, ex)
    try:
        cost = ctx.dtw_batch(A, B, it)
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    want = oracle.dtw_batch(A, B, it)
    bad = np.nonzero(bits(cost) != bits(want))[0]
    assert len(bad) == 0, [(shapes[k], float(cost[k]), float(want[k])) for k in bad[:8]]


def test_dtw_wave_classes(ctx, oracle):
    """Every rows-per-lane class of the wavefront kernel, band heights right at the class limits, n == m and n >> m, R' odd/even."""
    import rawalign_b200 as ra
    rng = np.random.default_rng(32)
    shapes = []
    for H in (1, 2, 4, 8, 16, 32):
        lim = 31 * H                                        # tallest band the class takes: 2R'+2 <= 31H
        for R in {max(1, (lim - 2) // 2), max(1, (lim - 2) // 2 - 1), max(1, lim // 4)}:
            shapes += [(6 * R + 40, 6 * R + 40, R), (6 * R + 60, 5 * R + 30, R), (8 * R + 100, 3 * R + 7, R)]
        shapes += [(lim + 50, lim, -1), (lim, lim + 13, -1), (2 * lim, max(1, lim - 1), -1)]
    shapes += [(40, 33, 1), (64, 64, 1), (65, 1, 3), (1, 77, 2), (33, 32, -1), (2000, 31, -1), (31, 2000, 5)]
    # one CTA per problem (256 threads, mailbox between warps): bands up to 255*H rows, H = 8, 16, 32
    shapes += [(3100, 3000, 1000), (2600, 2300, 1015), (6100, 6000, 2000), (5000, 9000, 2030), (12100, 12000, 4000), (9000, 8990, 4070),
               (1500, 1200, -1), (1100, 2039, -1), (3000, 2500, -1), (6000, 5000, -1), (8160, 8200, -1)]
    a_lens = np.array([s[0] for s in shapes]); b_lens = np.array([s[1] for s in shapes]); R = np.array([s[2] for s in shapes])
    ex = rng.integers(0, 2, len(shapes))
    A = rng.uniform(-2.5, 2.5, int(a_lens.sum())).astype(np.float32); B = rng.uniform(-2.5, 2.5, int(b_lens.sum())).astype(np.float32)
    it = _items(ra, a_lens, b_lens, R, ex)
    cost = ctx.dtw_batch(A, B, it)
    want = oracle.dtw_batch(A, B, it)
    bad = np.nonzero(bits(cost) != bits(want))[0]
    assert len(bad) == 0, [(shapes[k], float(cost[k]), float(want[k])) for k in bad[:8]]


def test_dtw_sparse_like_stream(ctx, oracle):
    """1e5 tiny anchor-to-anchor problems (a_len, b_len ~ 2+Geom(0.2), R = 1), overlapping spans."""
    import rawalign_b200 as ra
    rng = np.random.default_rng(32)
    n = 100_000
    a_lens = 1 + rng.geometric(0.2, n); b_lens = 1 + rng.geometric(0.2, n)
    A = rng.normal(0, 1, 50_000).astype(np.float32); B = rng.normal(0, 1, 80_000).astype(np.float32)
    a_off = rng.integers(0, len(A) - a_lens.max(), n); b_off = rng.integers(0, len(B) - b_lens.max(), n)
    it = ra.make_items(a_off, b_off, a_lens, b_lens, np.maximum(1, (a_lens * 0.1).astype(int)), rng.integers(0, 2, n))
    cost = ctx.dtw_batch(A, B, it)
    assert np.array_equal(bits(cost), bits(oracle.dtw_batch(A, B, it)))


def test_dtw_properties_at_full_size(ctx):
    """Size-independent properties on a batch too large for the oracle: identical spans cost 0;
    a band that covers the matrix equals the full matrix; cost is symmetric in (a, b)."""
    import rawalign_b200 as ra
    rng = np.random.default_rng(33)
    n = 4000
    a_lens = rng.integers(300, 900, n); b_lens = rng.integers(300, 900, n)
    A = rng.normal(0, 1, int(a_lens.sum())).astype(np.float32); B = rng.normal(0, 1, int(b_lens.sum())).astype(np.float32)
    it = _items(ra, a_lens, b_lens, 2 * np.maximum(a_lens, b_lens), np.zeros(n, int))
    wide = ctx.dtw_batch(A, B, it)
    it["band_radius"] = -1
    full = ctx.dtw_batch(A, B, it)
    assert np.array_equal(bits(wide), bits(full))
    sw = it.copy(); sw["a_off"], sw["b_off"] = it["b_off"], it["a_off"]; sw["a_len"], sw["b_len"] = it["b_len"], it["a_len"]
    assert np.array_equal(bits(ctx.dtw_batch(B, A, sw)), bits(full))
    same = ra.make_items(it["a_off"], it["a_off"], a_lens, a_lens, np.maximum(1, (a_lens * 0.1).astype(int)), np.zeros(n, int))
    assert np.all(ctx.dtw_batch(A, A, same) == 0.0)


def test_dtw_rejects_bad_items(ctx):
    import rawalign_b200 as ra
    A = np.zeros(10, np.float32)
    with pytest.raises(ra.RawAlignError):
        ctx.dtw_batch(A, A, ra.make_items([0], [0], [0], [5], [1], [0]))
    with pytest.raises(ra.RawAlignError):
        ctx.dtw_batch(A, A, ra.make_items([8], [0], [5], [5], [1], [0]))
    with pytest.raises(ValueError):
        ctx.DTW_global_slantedbanded_antidiagonalwise(A, A, -1)


def test_reference_named_single_calls(ctx, oracle):
    rng = np.random.default_rng(34)
    a = rng.normal(size=120).astype(np.float32); b = rng.normal(size=100).astype(np.float32)
    assert bits(ctx.DTW_global(a, b)) == bits(oracle.dtw_global(a, b))
    assert bits(ctx.DTW_global(a, b, True)) == bits(oracle.dtw_global(a, b, True))
    assert bits(ctx.DTW_global_slantedbanded_antidiagonalwise(a, b, 12)) == bits(oracle.dtw_banded(a, b, 12))
    c, pi, pj, pd = ctx.DTW_global_tb(a, b)
    c2, i2, j2, d2 = oracle.dtw_global_tb(a, b)
    assert bits(c) == bits(c2) and np.array_equal(pi, i2) and np.array_equal(pj, j2) and np.array_equal(bits(pd), bits(d2))


# ---- K4 -----------------------------------------------------------------------------------------------
def test_traceback_golden(ctx, dtw_golden):
    import rawalign_b200 as ra
    g = dtw_golden
    idx = g["tb_idx"]
    a_len = np.diff(g["a_off"])[idx]; b_len = np.diff(g["b_off"])[idx]
    it = ra.make_items(g["a_off"][:-1][idx], g["b_off"][:-1][idx], a_len, b_len, -np.ones(len(idx), int), g["exclude_last"][idx])
    cost, pi, pj, pd, off = ctx.dtw_tb_batch(g["a"], g["b"], it)
    assert np.array_equal(bits(cost), bits(g["tb_cost"]))
    assert np.array_equal(off.astype(np.int64), g["tb_off"])
    assert np.array_equal(pi, g["tb_i"]) and np.array_equal(pj, g["tb_j"])
    assert np.array_equal(bits(pd), bits(g["tb_diff"]))


def test_traceback_large_and_ties(ctx, oracle):
    import rawalign_b200 as ra
    rng = np.random.default_rng(35)
    # quantised values create many exact ties: the tie rule (diagonal wins, dtw.cpp:636-653) must match
    shapes = [(1500, 1400), (700, 2100), (33, 17), (1, 40), (40, 1), (1, 1), (260, 260)]
    a_lens = np.array([s[0] for s in shapes]); b_lens = np.array([s[1] for s in shapes])
    A = (rng.integers(-3, 4, int(a_lens.sum())) * 0.5).astype(np.float32); B = (rng.integers(-3, 4, int(b_lens.sum())) * 0.5).astype(np.float32)
    it = _items(ra, a_lens, b_lens, -np.ones(len(shapes), int), np.array([0, 0, 1, 0, 1, 0, 0]))
    cost, pi, pj, pd, off = ctx.dtw_tb_batch(A, B, it)
    for k in range(len(shapes)):
        a = A[int(it["a_off"][k]):int(it["a_off"][k]) + a_lens[k]]; b = B[int(it["b_off"][k]):int(it["b_off"][k]) + b_lens[k]]
        c2, i2, j2, d2 = oracle.dtw_global_tb(a, b, bool(it["exclude_last"][k]))
        s, e = int(off[k]), int(off[k + 1])
        assert bits(cost[k]) == bits(c2), k
        assert np.array_equal(pi[s:e], i2) and np.array_equal(pj[s:e], j2), k
        assert np.array_equal(bits(pd[s:e]), bits(d2)), k


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{"RA_FSM_SEG": "1"}, {"RA_FSM_SEG": "2"}, {"RA_FSM_SEG": "4", "RA_FSM_FORCE_REPLAY": "1"}])
def test_fsm_segment_modes(env):
    """The state machine's segment count and its replay path are fixed per process (environment), so the event parity
    tests are re-run in a child process for: the sequential form (1 segment), 2 segments, and 4 segments with every split
    chunk forced through the sequential replay."""
    import os, subprocess, sys
    e = dict(os.environ); e.update(env)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_parity.py"), "-q", "-x", "-m", "gpu",
                        "-k", "events and not segment_modes", "-p", "no:cacheprovider"], env=e, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


# ---- chain level --------------------------------------------------------------------------------------
def _workload(n_reads=40, seed=41):
    from workload import synth
    rng = np.random.default_rng(seed)
    n_ev = rng.integers(100, 2500, n_reads)
    ev_off = np.concatenate(([0], np.cumsum(n_ev)))
    read = rng.normal(0, 1, int(ev_off[-1])).astype(np.float32)
    ref = rng.normal(0, 1, 200_000).astype(np.float32)
    anchors, chains = synth.synth_chains(n_ev, ev_off[:-1], len(ref), seed=seed + 1, chains_per_read=4)
    # make the first chain of each read a true match so scores pass dtw_min_score and the running threshold bites
    for g in np.unique(chains["group"]):
        ch = chains[chains["group"] == g][0]
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        rd = read[int(ch["read_off"]):]
        for p in range(len(a) - 1, 0, -1):                   # anchor-to-anchor: the reference span is the read span, resampled
            qs, qe = int(a["query_position"][p]), int(a["query_position"][p - 1])
            ts, te = int(a["target_position"][p]), int(a["target_position"][p - 1])
            seg = rd[qs:qe + 1]
            ref[ts:te + 1] = np.interp(np.linspace(0, len(seg) - 1, te - ts + 1), np.arange(len(seg)), seg) + rng.normal(0, 0.05, te - ts + 1)
    return read, ref.astype(np.float32), anchors, chains


@pytest.mark.parametrize("border,fill", [(1, 1), (1, 0), (0, 1), (0, 0)])
def test_align_chains_batch_vs_oracle(ctx, oracle, border, fill):
    import rawalign_b200 as ra
    from oracle import DtwOpt
    read, ref, anchors, chains = _workload()
    opt = ra.ri_mapopt_init(); opt.dtw_border_constraint = border; opt.dtw_fill_method = fill
    ctx.set_opt(opt)
    try:
        score, keep, issued, useful = ctx.align_chains_batch(read, ref, anchors, chains, min_score_mode=1)
    finally:
        ctx.set_opt(ra.ri_mapopt_init())
    oo = DtwOpt(border, fill, 0.10, 0.4, 20.0)
    best, grp, useful_o, dropped = 0.0, None, 0, 0
    for k, ch in enumerate(chains):
        if grp != ch["group"]:
            best, grp = 0.0, ch["group"]
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        s, cells = oracle.align_chain(a, read[int(ch["read_off"]):], ref, oo, min_score=best, score_fma=1, return_cells=True)
        useful_o += cells
        assert bits(score[k]) == bits(s), (k, score[k], s)
        assert bool(keep[k]) == bool(s >= np.float32(20.0))
        dropped += s == np.float32(-1e10)
        if s >= np.float32(20.0) and s > best:
            best = float(s)
    assert useful == useful_o and issued >= useful
    assert keep.sum() >= len(np.unique(chains["group"])) // 2
    if border == 1:
        assert dropped > 0, "the early-termination replay was never exercised"


@pytest.mark.parametrize("legacy", [0, 1])       # 0: warp tiles (k_sp_parts2 + k_sp_rest), 1: CTA tiles (k_sp_parts)
@pytest.mark.parametrize("gap_p,frac", [(0.2, 0.10), (0.05, 0.10), (0.02, 0.25), (0.004, 0.10)])
def test_sparse_chain_part_shapes_vs_oracle(oracle, legacy, gap_p, frac):
    """Sparse-border chains whose parts are tiny (R' = 1 / 2: the register evaluator), medium (other narrow parts: full
    rounds in the tile and the deferred remainder), wide, and larger than a warp's staging window (gap_p 0.004: parts of
    hundreds of events, some beyond 1024) -- every class of k_sp_parts2, bit for bit against the oracle."""
    import rawalign_b200 as ra
    from workload import synth
    from oracle import DtwOpt
    rng = np.random.default_rng(int(1000 * gap_p) + 7)
    n_reads = 12
    n_ev = rng.integers(600, 5000, n_reads)
    ev_off = np.concatenate(([0], np.cumsum(n_ev)))
    read = rng.normal(0, 1, int(ev_off[-1])).astype(np.float32)
    ref = rng.normal(0, 1, 120_000).astype(np.float32)
    anchors, chains = synth.synth_chains(n_ev, ev_off[:-1], len(ref), seed=11, chains_per_read=3, gap_p=gap_p)
    opt = ra.ri_mapopt_init(); opt.dtw_band_radius_frac = frac; opt.reserved = 8 * legacy; opt.dtw_min_score = -1e9
    c = ra.Context(0, opt)
    try:
        ms = np.full(len(chains), -1e10, np.float32)                       # no early termination: every part is evaluated
        score, keep, issued, useful = c.align_chains_batch(read, ref, anchors, chains, min_score_mode=0, min_score=ms)
    finally:
        c.close()
    oo = DtwOpt(1, 1, frac, 0.4, -1e9)
    tot = 0
    for k, ch in enumerate(chains):
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        sc, cells = oracle.align_chain(a, read[int(ch["read_off"]):], ref, oo, min_score=-1e10, score_fma=1, return_cells=True)
        tot += cells
        assert bits(score[k]) == bits(sc), (k, score[k], sc)
    assert useful == tot and issued == tot


def test_align_chain_cigar_vs_oracle(ctx, oracle):
    import rawalign_b200 as ra
    read, ref, anchors, chains = _workload(n_reads=6, seed=43)
    chains = chains[::4]
    for border in (1, 0):
        opt = ra.ri_mapopt_init(); opt.dtw_border_constraint = border; opt.dtw_fill_method = 0
        ctx.set_opt(opt)
        try:
            score, cost, pi, pj, pd, off = ctx.align_chains_cigar_batch(read, ref, anchors, chains)
        finally:
            ctx.set_opt(ra.ri_mapopt_init())
        for k, ch in enumerate(chains):
            a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
            rd = read[int(ch["read_off"]):]
            wi, wj, wd, c_tot, n_al = [], [], [], np.float32(0), 0
            if border == 1:                                  # rmap.cpp:282-292
                parts = len(a) - 1
                for p in range(parts):
                    s, e = a[parts - p], a[parts - p - 1]
                    c, i2, j2, d2 = oracle.dtw_global_tb(rd[s["query_position"]:e["query_position"] + 1], ref[s["target_position"]:e["target_position"] + 1])
                    wi.append(i2.astype(np.uint64) + int(s["query_position"])); wj.append(j2.astype(np.uint64) + int(s["target_position"])); wd.append(d2)
                    c_tot = np.float32(c_tot + c); n_al += int(e["query_position"]) - int(s["query_position"]) + 1
            else:                                            # rmap.cpp:218-235 (offset applied to back() only, n times)
                s, e = a[-1], a[0]
                c, i2, j2, d2 = oracle.dtw_global_tb(rd[s["query_position"]:e["query_position"] + 1], ref[s["target_position"]:e["target_position"] + 1])
                i2 = i2.astype(np.uint64); j2 = j2.astype(np.uint64)
                i2[-1] += len(i2) * int(s["query_position"]); j2[-1] += len(j2) * int(s["target_position"])
                wi, wj, wd = [i2], [j2], [d2]; c_tot = c; n_al = int(e["query_position"]) - int(s["query_position"]) + 1
            wi, wj, wd = np.concatenate(wi), np.concatenate(wj), np.concatenate(wd)
            s0, e0 = int(off[k]), int(off[k + 1])
            assert np.array_equal(pi[s0:e0], wi) and np.array_equal(pj[s0:e0], wj) and np.array_equal(bits(pd[s0:e0]), bits(wd))
            assert bits(cost[k]) == bits(c_tot)
            # fused n*bonus - cost (rmap.cpp:306 as built): the float64 product of two float32 is exact
            want = np.float32(np.float64(np.float32(n_al)) * np.float64(np.float32(0.4)) - np.float64(c_tot))
            assert bits(score[k]) == bits(want)
    opt = ra.ri_mapopt_init(); opt.dtw_border_constraint = 0; opt.dtw_fill_method = 1
    ctx.set_opt(opt)
    try:
        with pytest.raises(ra.RawAlignError) as e:
            ctx.align_chains_cigar_batch(read, ref, anchors, chains)
        assert e.value.code == -5
    finally:
        ctx.set_opt(ra.ri_mapopt_init())


# ---- raw-signal decode + sliced/overlapped entry points -------------------------------------------------
def _oracle_raw_pipeline(oracle, raw, raw_off, raw_len, offset, scale, chunk_size, max_chunks):
    sigs, coff, clen, first, lsig = [], [], [], [0], []
    pos = 0
    for r in range(len(raw_len)):
        a = int(raw_off[r])
        d = oracle.decode_raw(raw[a:a + int(raw_len[r])], offset[r], scale[r])
        lsig.append(len(d))
        c = 0; cs = 0
        while cs < len(d) and c < max_chunks:                      # rmap.cpp:684-689
            coff.append(pos + cs); clen.append(min(chunk_size, len(d) - cs)); cs += chunk_size; c += 1
        first.append(len(coff))
        sigs.append(d); pos += len(d)
    sig = np.concatenate(sigs) if sigs else np.zeros(0, np.float32)
    ev, off = oracle.detect_events_batch(sig, np.array(coff, np.uint64), np.array(clen, np.uint32))
    return ev, off, np.array(lsig, np.uint32), np.array(first, np.uint32)


@pytest.mark.parametrize("align", [8, 1])
def test_events_raw_batch_vs_oracle(ctx, oracle, align):
    from workload import synth
    rng = np.random.default_rng(77)
    ref = synth.random_reference(30_000, seed=1)
    rd = synth.simulate_reads(ref, 24, seed=12, max_chunks=3)
    # re-lay the reads with the requested alignment, ragged lengths, and samples outside (30, 200) pA to exercise the filter
    raws, offs, lens = [], [], []
    pos = 5 if align == 1 else 0
    for r in range(24):
        a = int(rd.read_off[r]); n = int(rd.read_len[r]) - int(rng.integers(0, 50))
        x = rd.raw[a:a + n].copy()
        bad = rng.random(n) < 0.03
        x[bad] = rng.choice(np.array([0, 40, 100, 1200, 3000, -500], np.int16), int(bad.sum()))
        if r == 3: x = x[:7]                                       # shorter than the detector windows
        if r == 5: x[:] = 3000                                     # everything filtered out -> no chunk at all
        raws.append(x); offs.append(pos); lens.append(len(x)); pos += len(x); pos = (pos + align - 1) // align * align
    raw = np.zeros(pos + 16, np.int16)
    for x, o in zip(raws, offs):
        raw[o:o + len(x)] = x
    raw_off = np.array(offs, np.uint64); raw_len = np.array(lens, np.uint32)
    offset = rd.dac_offset[:24].copy(); scale = rd.dac_scale[:24].copy()
    offset[7] = np.float32(-3.5); scale[9] = np.float32(0.21)       # per-read channel attributes differ
    for max_chunks in (2, 30):
        ev, off, lsig, first = ctx.detect_events_raw_batch(raw, raw_off, raw_len, offset, scale, 4000, max_chunks)
        ev_o, off_o, lsig_o, first_o = _oracle_raw_pipeline(oracle, raw, raw_off, raw_len, offset, scale, 4000, max_chunks)
        assert np.array_equal(lsig, lsig_o), "filtered lengths differ"
        assert np.array_equal(first, first_o), "chunk tables differ"
        assert np.array_equal(off, off_o), "event counts differ"
        assert np.array_equal(bits(ev), bits(ev_o)), "event values differ"


def test_events_sliced_copies_large_batch(ctx, oracle):
    """Enough signal (> 64 MB) for the host->device copy to be cut into several slices that overlap K1."""
    from workload import synth
    ref = synth.random_reference(60_000, seed=1)
    rd = synth.simulate_reads(ref, 640, seed=21, max_chunks=8)
    coff, clen, _ = synth.chunk_table(rd.read_off, rd.read_len, max_chunks=8)
    assert int(clen.sum()) > (16 << 20)
    ev, off = ctx.detect_events_batch(rd.sig, coff, clen)
    ev_o, off_o = oracle.detect_events_batch(rd.sig, coff, clen)
    assert np.array_equal(off, off_o) and np.array_equal(bits(ev), bits(ev_o))
    ev2, off2, lsig, first = ctx.detect_events_raw_batch(rd.raw, rd.read_off[:-1], rd.read_len, rd.dac_offset, rd.dac_scale, 4000, 8)
    assert np.array_equal(lsig, rd.read_len) and int(first[-1]) == len(coff)
    assert np.array_equal(off2, off_o) and np.array_equal(bits(ev2), bits(ev_o))


# ---- sliced chain upload and anchor validation ------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("slice_anchors", [200, 1500, 1 << 30])
def test_align_chains_sliced_upload_vs_oracle(oracle, slice_anchors):
    """The sliced upload path of ra_align_chains_batch (h2d stream, per-slice group cuts and pipelines) with 8 slices, a few
    slices (some empty) and the one-shot path give the oracle's scores bit for bit."""
    import rawalign_b200 as ra
    from oracle import DtwOpt
    read, ref, anchors, chains = _workload(n_reads=30, seed=91)
    oo = DtwOpt(1, 1, 0.10, 0.4, 20.0)
    want = np.zeros(len(chains), np.float32)
    best, grp = 0.0, None
    for k, ch in enumerate(chains):
        if grp != ch["group"]:
            best, grp = 0.0, ch["group"]
        a = anchors[int(ch["anchor_off"]):int(ch["anchor_off"]) + int(ch["n_anchors"])]
        s = oracle.align_chain(a, read[int(ch["read_off"]):], ref, oo, min_score=best, score_fma=1)
        want[k] = s
        if s >= np.float32(20.0) and s > best:
            best = float(s)
    with ra.Context(0) as c:
        c.set_tuning("chain_slice_min_anchors", slice_anchors)
        score, keep, issued, useful = c.align_chains_batch(read, ref, anchors, chains, min_score_mode=1)
        assert np.array_equal(bits(score), bits(want))
        assert np.array_equal(keep.astype(bool), want >= np.float32(20.0))
        # a second call on the same context (counters and staging reused)
        score2, _, _, _ = c.align_chains_batch(read, ref, anchors, chains, min_score_mode=1, count_cells=False)
        assert np.array_equal(bits(score2), bits(want))


@pytest.mark.gpu
@pytest.mark.parametrize("slice_anchors", [300, 1 << 30])
def test_align_chains_rejects_malformed_anchors(slice_anchors):
    """A non-monotonic interior anchor or a chain end outside the arenas is RA_EINVAL, not an out-of-bounds access."""
    import rawalign_b200 as ra
    read, ref, anchors, chains = _workload(n_reads=12, seed=92)
    with ra.Context(0) as c:
        c.set_tuning("chain_slice_min_anchors", slice_anchors)
        bad = anchors.copy()
        ch = chains[len(chains) // 2]
        bad["query_position"][int(ch["anchor_off"]) + int(ch["n_anchors"]) // 2] += 100000       # interior anchor jumps ahead of its neighbours
        with pytest.raises(ra.RawAlignError) as e:
            c.align_chains_batch(read, ref, bad, chains, min_score_mode=1)
        assert e.value.code == -3
        bad = anchors.copy()
        bad["target_position"][int(ch["anchor_off"]) + 1] = 0xFFFFFFF0
        with pytest.raises(ra.RawAlignError) as e:
            c.align_chains_batch(read, ref, bad, chains, min_score_mode=1)
        assert e.value.code == -3
        score, _, _, _ = c.align_chains_batch(read, ref, anchors, chains, min_score_mode=1)       # the context still works
        assert np.isfinite(score).all()
