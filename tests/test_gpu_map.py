"""-m gpu: the map loop on the device (ra_index_load_*, ra_map_batch: seeding, index lookup, anchor sort, chaining DP,
DTW evaluation, primary chains, MAPQ, chunk loop, PAF fields) through the C ABI against
  * the PAF lines the reference's own map_worker_for printed (tests/golden/map_cases.npz, oracle/gen_golden_map.py), and
  * the pinned CPU restatement (oracle/ra_oracle_map.c) on fresh random data, field by field."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mg():
    g = np.load(os.path.join(GOLDEN, "map_cases.npz"))
    return {k: g[k] for k in g.files}


@pytest.fixture(scope="module")
def flat(mg):
    from oracle import build_flat_index
    return {p: build_flat_index([mg["seq0"], mg["seq1"]], mg["pore_vals"], p) for p in ("sensitive", "viral", "fast")}


def _opt(kw):
    import rawalign_b200 as ra
    return ra.ri_mapopt_full(**kw)


def _dac(n):
    from workload import synth
    return (np.full(n, np.float32(synth.DAC_OFFSET)), np.full(n, np.float32(np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))))


@pytest.mark.parametrize("ci", range(6))
def test_map_batch_paf_vs_reference_golden(mg, flat, ci):
    import rawalign_b200 as ra
    from oracle.gen_golden_map import CONFIGS
    preset, kw = CONFIGS[ci]
    off, sc = _dac(len(mg["raw_len"]))
    with ra.Context(0) as ctx, ra.Index.load_flat(ctx, flat[preset]) as ix:
        rec, st = ix.map_reads(_opt(kw), raw=mg["raw"], raw_off=mg["raw_off"], raw_len=mg["raw_len"], offset=off, scale=sc)
        got = ix.paf_lines(rec)
        rec2, st2 = ix.map_reads(_opt(kw), resident_reads=len(mg["raw_len"]), count_cells=True)      # device-resident re-run
    want = str(mg["paf"][ci]).split("\n")
    assert len(got) == len(want)
    for r, (a, b) in enumerate(zip(got, want)):
        assert a == b, (r, a, b)
    assert 1 <= st["rounds"] <= 30 and st["launches"] > 0
    assert rec2.tobytes() == rec.tobytes() and st2["chunks"] == st["chunks"]
    assert (st2["cells_useful"] > 0) == bool(kw.get("flag", 0) & 2)


def test_index_file_loader_and_float_input(mg, flat, tmp_path):
    """ra_index_load_file on a file in the reference's format; float pA input instead of DAC samples; same PAF."""
    import rawalign_b200 as ra
    from oracle import write_index_file
    from oracle.gen_golden_map import CONFIGS
    from workload import synth
    p = str(tmp_path / "ref.ind")
    write_index_file(p, flat["sensitive"])
    sig = ((mg["raw"].astype(np.float32) + np.float32(synth.DAC_OFFSET)) * (np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))).astype(np.float32)
    with ra.Context(0) as ctx, ra.Index.load_file(ctx, p) as ix:
        assert ix.n_seq == 2 and ix.e == 6 and ix.names == ["ref0", "ref1"]
        rec, _ = ix.map_reads(_opt(CONFIGS[0][1]), sig=sig, sig_off=mg["raw_off"], sig_len=mg["raw_len"])
        got = ix.paf_lines(rec)
    assert got == str(mg["paf"][0]).split("\n")


@pytest.mark.parametrize("preset,kw,n_seq", [("sensitive", dict(flag=2), 1), ("fast", dict(flag=2, dtw_border_constraint=0, dtw_fill_method=1), 3),
                                             ("viral", dict(flag=0, max_num_chunk=6), 2), ("sensitive", dict(flag=2, min_events=600, num_best_chains=5), 1)])
def test_map_batch_vs_oracle_random(preset, kw, n_seq):
    """Fresh data: clean, noisy and junk reads over 1-3 sequences; every PAF field equals the pinned restatement's."""
    import rawalign_b200 as ra
    from oracle import MapOracle, build_flat_index
    from workload import synth
    rng = np.random.default_rng(1000 + n_seq)
    seqs = [synth.random_reference(int(n), seed=50 + i) for i, n in enumerate([400_000, 120_000, 60_000][:n_seq])]
    tab = build_flat_index(seqs, synth.pore_model_as_loaded(), preset)
    rd = synth.simulate_reads(seqs[0], 96, seed=9, max_chunks=30)
    sig = rd.sig.copy()
    for r in range(0, 96, 8):                                      # junk
        a = int(rd.read_off[r]); n = int(rd.read_len[r]); sig[a:a + n] = np.clip(np.repeat(rng.normal(90, 13, n // 8 + 1), 8)[:n] + rng.normal(0, 2, n), 31, 199)
    for r in range(3, 96, 5):                                      # noisy
        a = int(rd.read_off[r]); n = int(rd.read_len[r]); sig[a:a + n] = np.clip(sig[a:a + n] + rng.normal(0, 5.0, n), 31, 199)
    sig = sig.astype(np.float32)
    mo = MapOracle(tab)
    want = mo.map_reads(mo.mapopt(**kw), sig, rd.read_off[:-1], rd.read_len)
    with ra.Context(0) as ctx, ra.Index.load_flat(ctx, tab) as ix:
        rec, st = ix.map_reads(_opt(kw), sig=sig, sig_off=rd.read_off[:-1], sig_len=rd.read_len)
        got_lines = ix.paf_lines(rec)
    for f in ("mapped", "ref_id", "read_length", "read_start_position", "read_end_position", "fragment_start_position", "fragment_length", "mapq",
              "rev", "ci", "sl", "cm", "nc", "n_events"):
        assert np.array_equal(rec[f], want[f]), (f, np.nonzero(rec[f] != want[f])[0][:8])
    for f in ("s1", "s2", "sm", "at", "aq"):
        assert np.array_equal(bits(rec[f]), bits(want[f])), f
    m = want["mapped"].astype(bool)
    assert np.array_equal(bits(rec["alns"][m]), bits(want["alns"][m]))
    assert got_lines == mo.paf_lines(want)
    if kw.get("min_events", 50) == 50:
        assert want["mapped"].sum() >= 40 and (want["ci"] > 1).sum() >= 5


@pytest.mark.parametrize("ci,flags", [(0, ["--dtw-evaluate-chains"]), (1, []),
                                      (5, ["--dtw-evaluate-chains", "--dtw-border-constraint", "global", "--dtw-fill-method", "banded=0.1"])])
def test_compiled_cpp_consumer_prints_the_reference_paf(mg, flat, tmp_path, ci, flags):
    """integration/rawalign_b200_map.cpp: a C++ program compiled against include/rawalign_b200.h and linked with the
    library (no Python, no ctypes) maps the fixture reads from the reference-format index file and prints PAF lines that equal
    the reference's map_worker_for output (minus mt:f:)."""
    import re
    import subprocess
    from oracle import write_index_file
    from oracle.gen_golden_map import CONFIGS
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "integration"), "-B", "rawalign_b200_map"], stdout=subprocess.DEVNULL)
    preset, kw = CONFIGS[ci]
    ind = str(tmp_path / "ref.ind"); rb = str(tmp_path / "reads.bin")
    write_index_file(ind, flat[preset])
    n = len(mg["raw_len"]); off, sc = _dac(n)
    with open(rb, "wb") as f:
        f.write(np.uint32(n).tobytes()); f.write(mg["raw_off"].astype(np.uint64).tobytes()); f.write(mg["raw_len"].astype(np.uint32).tobytes())
        f.write(off.astype(np.float32).tobytes()); f.write(sc.astype(np.float32).tobytes())
        f.write(np.uint64(len(mg["raw"])).tobytes()); f.write(mg["raw"].astype(np.int16).tobytes())
    r = subprocess.run([os.path.join(root, "integration", "rawalign_b200_map"), "--batch", "16"] + flags + [ind, rb], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = [re.sub(r"mt:f:[0-9.eE+-]+\t", "", l) for l in r.stdout.splitlines()]
    assert got == str(mg["paf"][ci]).split("\n")


def test_map_batch_cigar_vs_reference_golden(mg, flat):
    """--dtw-output-cigar through the map loop: alns:f: and every (i,j,diff) triple of aln:s: equal the reference's."""
    import rawalign_b200 as ra
    off, sc = _dac(len(mg["raw_len"]))
    n = len(mg["raw_len"])
    with ra.Context(0) as ctx, ra.Index.load_flat(ctx, flat["sensitive"]) as ix:
        rec, st = ix.map_reads(_opt(dict(flag=ra.RI_M_DTW_EVALUATE_CHAINS | ra.RI_M_DTW_OUTPUT_CIGAR)), raw=mg["raw"], raw_off=mg["raw_off"], raw_len=mg["raw_len"], offset=off, scale=sc)
        cig = ix.last_cigar(n)
        got = ix.paf_lines(rec, cigar=cig)
    want = str(mg["paf_cigar"]).split("\n")
    assert len(got) == len(want)
    for r, (a, b) in enumerate(zip(got, want)):
        assert a == b, (r, a[:300], b[:300])
    assert int(cig[0][-1]) > 10000


def test_bitonic_sort_fallback_gives_the_same_paf():
    """The anchor sort has two forms (bucket sort in shared memory; the bitonic network for segments beyond it, or when the
    index has too many sequences for the bin table).  RA_MAP_BITONIC_ONLY=1 sends every segment through the network: the golden
    PAF tests are re-run that way in a child process."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ); e["RA_MAP_BITONIC_ONLY"] = "1"
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_map.py"), "-q", "-x", "-m", "gpu", "-k", "paf_vs_reference_golden",
                        "-p", "no:cacheprovider"], env=e, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_index_replica_by_copy(mg, flat, tmp_path):
    """ra_index_alloc + ra_index_tables: a replica allocated from rank 0's geometry and filled by device-to-device copies of the
    three tables (what the NCCL broadcast does across GPUs) maps to the same PAF."""
    import torch
    import rawalign_b200 as ra
    from oracle.gen_golden_map import CONFIGS
    off, sc = _dac(len(mg["raw_len"]))
    with ra.Context(0) as c0, ra.Context(0) as c1, ra.Index.load_flat(c0, flat["sensitive"]) as ix0:
        info = ix0.counts(); tabs = ix0.tables()
        rep = ra.Index.alloc_like(c1, info, (tabs[0][1] // 16 - 16) // 2, tabs[1][1] // 8, ix0.seq_len, ix0.names)
        try:
            assert [b for _, b in rep.tables()] == [b for _, b in tabs]
            for (src, n), (dst, _) in zip(tabs, rep.tables()):
                class _V:
                    def __init__(self, p):
                        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (p, False), "version": 2}
                torch.as_tensor(_V(dst), device="cuda:0").copy_(torch.as_tensor(_V(src), device="cuda:0"))
            torch.cuda.synchronize()
            rec, _ = rep.map_reads(_opt(CONFIGS[0][1]), raw=mg["raw"], raw_off=mg["raw_off"], raw_len=mg["raw_len"], offset=off, scale=sc)
            assert rep.paf_lines(rec) == str(mg["paf"][0]).split("\n")
        finally:
            rep.close()


def _dev_f32(ptr, n):
    import torch

    class _V:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}
    return torch.as_tensor(_V(), device="cuda:0").cpu().numpy()


@pytest.mark.parametrize("preset,lens,contract", [("sensitive", [300_000], 1), ("fast", [150_000, 9, 40_000, 777], 1), ("viral", [29_903, 5_000], 0)])
def test_index_build_on_device_equals_the_reference_table(preset, lens, contract):
    """f4: ra_index_build (ri_idx_gen on the device) -- the exported seed table, the expected signals in the mirror and the PAF
    mapped through the built index equal those of the pinned restatement of rawindex.cpp / rsig.cpp (which test_map_oracle pins
    against index files the reference wrote).  Ambiguous bases, a sequence of exactly k bases, both contraction modes."""
    import rawalign_b200 as ra
    from oracle import PRESETS, build_flat_index
    from oracle.gen_golden_map import CONFIGS
    from workload import synth
    rng = np.random.default_rng(77)
    seqs = [synth.random_reference(int(n), seed=300 + i) for i, n in enumerate(lens)]
    for s in seqs:
        if len(s) > 1000:
            s[rng.integers(0, len(s), 40)] = 4                       # N bases (seq_nt4_table -> 4)
            s[500:530] = 4
    pore = synth.pore_model_as_loaded()
    tab = build_flat_index(seqs, pore, preset, contract=contract)
    e, q, lq, w, n = PRESETS[preset]
    rd = synth.simulate_reads(np.where(seqs[0] < 4, seqs[0], 0).astype(np.uint8), 48, seed=21, max_chunks=6)
    sig = rd.sig.astype(np.float32)
    kw = CONFIGS[0][1]
    with ra.Context(0) as ctx:
        o = ctx.opt; o.fp_contract = contract; ctx.set_opt(o)
        with ra.Index.build(ctx, seqs, pore, e=e, q=q, lq=lq, k=6) as ix:      # (one index per context at a time: the mirror is the context's)
            got = ix.export()
            assert np.array_equal(got["keys"], tab["keys"]) and np.array_equal(got["key_off"], tab["key_off"]) and np.array_equal(got["vals"], tab["vals"])
            assert ix.counts()[6:] == [len(seqs), len(tab["keys"]), len(tab["vals"])]
            ptr, nbytes = ix.tables()[2]
            mirror = _dev_f32(ptr, nbytes // 4)
            want = np.concatenate([np.concatenate((f, r)) for f, r in zip(tab["fwd"], tab["rev"])])
            assert mirror.shape == want.shape and np.array_equal(bits(mirror), bits(want))
            rec, _ = ix.map_reads(_opt(kw), sig=sig, sig_off=rd.read_off[:-1], sig_len=rd.read_len)
            got_paf = ix.paf_lines(rec)
        with ra.Index.load_flat(ctx, tab) as ref_ix:
            rec2, _ = ref_ix.map_reads(_opt(kw), sig=sig, sig_off=rd.read_off[:-1], sig_len=rd.read_len)
            assert got_paf == ref_ix.paf_lines(rec2)
            assert rec2["mapped"].sum() >= 20


def test_index_build_rejects_bad_input():
    import rawalign_b200 as ra
    from workload import synth
    pore = synth.pore_model_as_loaded()
    with ra.Context(0) as ctx:
        with pytest.raises(ra.RawAlignError):
            ra.Index.build(ctx, [np.zeros(5, np.uint8)], pore, e=6, q=9, lq=3, k=6)          # shorter than k
        with pytest.raises(ra.RawAlignError):
            ra.Index.build(ctx, [np.zeros(500, np.uint8)], pore, e=12, q=9, lq=3, k=6)        # e out of range
        with pytest.raises(ValueError):
            ra.Index.build(ctx, [np.zeros(500, np.uint8)], pore[:100], e=6, q=9, lq=3, k=6)
        with ra.Index.build(ctx, [np.arange(6, dtype=np.uint8) % 4, np.arange(40, dtype=np.uint8) % 4], pore, e=6, q=9, lq=3, k=6) as ix:
            assert ix.counts()[6] == 2                                                        # one k-mer: no seeds, no crash


def test_index_clone_for_a_second_context(mg, flat):
    """ra_index_clone: the batches in flight on one GPU share one parse of the index file."""
    import rawalign_b200 as ra
    from oracle.gen_golden_map import CONFIGS
    off, sc = _dac(len(mg["raw_len"]))
    with ra.Context(0) as c0, ra.Context(0) as c1, ra.Index.load_flat(c0, flat["sensitive"]) as ix0, ix0.clone(c1) as ix1:
        assert ix1.counts() == ix0.counts() and ix1.names == ix0.names
        rec, _ = ix1.map_reads(_opt(CONFIGS[0][1]), raw=mg["raw"], raw_off=mg["raw_off"], raw_len=mg["raw_len"], offset=off, scale=sc)
        assert ix1.paf_lines(rec) == str(mg["paf"][0]).split("\n")
        a, b = ix0.export(), ix1.export()
        assert all(np.array_equal(a[k], b[k]) for k in a)


def test_float_form_uploads_per_round_from_pinned_and_pageable_memory(mg, flat):
    """The float form moves only the chunk each active read is at: straight from page-locked caller memory (mapped reads by the
    GPU, any alignment), through a staging buffer from pageable memory; "map_sig_whole" uploads everything and leaves the batch
    resident.  Same records every way, and the bytes moved are the chunks consumed."""
    import torch
    import rawalign_b200 as ra
    from oracle.gen_golden_map import CONFIGS
    from workload import synth
    sig = ((mg["raw"].astype(np.float32) + np.float32(synth.DAC_OFFSET)) * (np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))).astype(np.float32)
    pinned = torch.from_numpy(sig).pin_memory().numpy()
    shifted = torch.from_numpy(np.concatenate((np.zeros(1, np.float32), sig))).pin_memory().numpy()      # every chunk starts 4 bytes off a 16-byte line
    n = len(mg["raw_len"])
    off, ln = mg["raw_off"], mg["raw_len"]
    opt = _opt(CONFIGS[0][1])
    with ra.Context(0) as ctx, ra.Index.load_flat(ctx, flat["sensitive"]) as ix:
        rec_a, st_a = ix.map_reads(opt, sig=sig, sig_off=off, sig_len=ln)
        rec_b, st_b = ix.map_reads(opt, sig=pinned, sig_off=off, sig_len=ln)
        rec_s, st_s = ix.map_reads(opt, sig=shifted, sig_off=off + np.uint64(1), sig_len=ln)
        with pytest.raises(ra.RawAlignError):
            ix.map_reads(opt, resident_reads=n)                     # nothing was left on the device
        ctx.set_tuning("map_sig_whole", 1)
        rec_c, st_c = ix.map_reads(opt, sig=pinned, sig_off=off, sig_len=ln)
        rec_d, _ = ix.map_reads(opt, resident_reads=n)
        got = ix.paf_lines(rec_b)
    assert got == str(mg["paf"][0]).split("\n")
    assert rec_a.tobytes() == rec_b.tobytes() == rec_s.tobytes() == rec_c.tobytes() == rec_d.tobytes()
    consumed = 4 * sum(int(min(4000, max(0, int(l) - 4000 * c))) for l, ci in zip(ln, rec_b["ci"]) for c in range(int(ci)))
    for st in (st_a, st_b, st_s):
        assert 0 <= st["h2d_bytes"] - consumed < 64 * n + 4096
    assert st_c["h2d_bytes"] >= int(ln.sum()) * 4 > st_b["h2d_bytes"]


def test_repetitive_reference_long_anchor_segments():
    """A tandem-repeat reference (500 copies of a 600-base unit) gives every chunk tens of thousands of anchors: the anchor sort's largest size class
    (binned keys staged in global memory, runs of bins finished in shared memory) and long chaining lists; all PAF fields equal
    the pinned restatement's."""
    import rawalign_b200 as ra
    from oracle import MapOracle, build_flat_index
    from workload import synth
    unit = synth.random_reference(600, seed=71)
    ref = np.concatenate([synth.random_reference(20_000, seed=72)] + [unit] * 500 + [synth.random_reference(20_000, seed=73)])
    tab = build_flat_index([ref], synth.pore_model_as_loaded(), "sensitive")
    rd = synth.simulate_reads(ref[20_000:20_000 + 300_000], 12, seed=5, max_chunks=3)
    sig = rd.sig.astype(np.float32)
    kw = dict(flag=2, max_num_chunk=3)
    mo = MapOracle(tab)
    want = mo.map_reads(mo.mapopt(**kw), sig, rd.read_off[:-1], rd.read_len)
    with ra.Context(0) as ctx, ra.Index.load_flat(ctx, tab) as ix:
        rec, st = ix.map_reads(_opt(kw), sig=sig, sig_off=rd.read_off[:-1], sig_len=rd.read_len)
        got_lines = ix.paf_lines(rec)
    assert st["anchors"] > 12 * 20480
    assert got_lines == mo.paf_lines(want)
    for f in ("s1", "s2", "sm", "at", "aq"):
        assert np.array_equal(bits(rec[f]), bits(want[f])), f


def test_index_build_on_device_equals_the_reference_indexer_output(tmp_path):
    """f4 pinned directly: the reference's unmodified indexer (oracle/_ref/rawalign -d) writes an index file for a FASTA; the table
    and signals ra_index_build makes from the same bases are identical to the parsed file."""
    import subprocess
    import rawalign_b200 as ra
    from oracle import PRESETS, parse_index_file
    from workload import synth
    from workload.mapbench import _ref_binary
    rb = _ref_binary()
    if rb is None:
        pytest.skip("oracle/_ref/rawalign was not built in this checkout")
    seqs = [synth.random_reference(n, seed=400 + i) for i, n in enumerate([250_000, 31_000, 4_321])]
    fa, model, ind = str(tmp_path / "r.fa"), str(tmp_path / "m.txt"), str(tmp_path / "r.ind")
    synth.write_fasta(fa, seqs); synth.write_pore_model(model)
    r = subprocess.run([rb, "-x", "fast", "-t", "4", "-p", model, "-d", ind, fa], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    tab = parse_index_file(ind)
    e, q, lq, w, n = PRESETS["fast"]
    assert (tab["e"], tab["q"], tab["lq"]) == (e, q, lq)
    with ra.Context(0) as ctx, ra.Index.build(ctx, seqs, synth.pore_model_as_loaded(), e=e, q=q, lq=lq, k=6) as ix:
        got = ix.export()
        ptr, nbytes = ix.tables()[2]
        mirror = _dev_f32(ptr, nbytes // 4)
    assert np.array_equal(got["keys"], tab["keys"]) and np.array_equal(got["key_off"], tab["key_off"]) and np.array_equal(got["vals"], tab["vals"])
    want = np.concatenate([np.concatenate((f, r_)) for f, r_ in zip(tab["fwd"], tab["rev"])])
    assert np.array_equal(bits(mirror), bits(want))


@pytest.mark.parametrize("border,fill", [(0, 0), (1, 0), (1, 1)])
def test_cigar_stage_vs_live_reference(tmp_path, border, fill):
    """--dtw-output-cigar under global border + full matrix (the last-element quirk of rmap.cpp:230-233), sparse + full and sparse +
    banded, against the reference's own map loop run live (oracle/_ref/libra_refmap.so), on an index built on the device."""
    import re
    import rawalign_b200 as ra
    from oracle import PRESETS, RefMap
    from workload import synth
    if not RefMap.available():
        pytest.skip("oracle/_ref/libra_refmap.so was not built in this checkout")
    seqs = [synth.random_reference(n, seed=500 + i) for i, n in enumerate([200_000, 50_000])]
    fa, model = str(tmp_path / "r.fa"), str(tmp_path / "m.txt")
    synth.write_fasta(fa, seqs); synth.write_pore_model(model)
    rd = synth.simulate_reads(seqs[0], 40, seed=31, max_chunks=4)
    sig = rd.sig.astype(np.float32)
    off, ln = rd.read_off[:-1].astype(np.uint64), rd.read_len.astype(np.uint32)
    kw = dict(flag=ra.RI_M_DTW_EVALUATE_CHAINS | ra.RI_M_DTW_OUTPUT_CIGAR, dtw_border_constraint=border, dtw_fill_method=fill)
    rm = RefMap()
    idx = rm.index_build(fa, model, "sensitive")
    paf, mapped, _ = rm.map_reads(idx, rm.mapopt(**kw), sig, off, ln, n_threads=4)
    rm.index_free(idx)
    want = [re.sub(r"mt:f:[0-9.eE+-]+\t", "", l) for l in paf]
    e, q, lq, w, n = PRESETS["sensitive"]
    with ra.Context(0) as ctx, ra.Index.build(ctx, seqs, synth.pore_model_as_loaded(), e=e, q=q, lq=lq, k=6) as ix:
        rec, _ = ix.map_reads(_opt(kw), sig=sig, sig_off=off, sig_len=ln)
        got = ix.paf_lines(rec, cigar=ix.last_cigar(len(ln)))
    assert mapped.sum() >= 20
    assert len(got) == len(want)
    for r, (a, b) in enumerate(zip(got, want)):
        assert a == b, (r, a[:300], b[:300])
