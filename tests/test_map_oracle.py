"""The map loop around the hot path (SURVEY.md rows f1 chaining, f2 sketch + index lookup, f4 reference -> expected signal,
and gen_chains / gen_primary_chains / comp_mapq / is_mapped_with_high_confidence / map_worker_for): the C restatement
oracle/ra_oracle_map.c against fixtures generated from the reference's compiled code (oracle/gen_golden_map.py), and live
against oracle/_ref/libra_refmap.so where that exists."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits
from oracle import ANCHOR_DTYPE, CHAIN_OUT_DTYPE, MapOracle, RefMap, build_flat_index
from oracle.gen_golden_map import CONFIGS, strip_mt, table_checksum

needs_refmap = pytest.mark.skipif(not RefMap.available(), reason="oracle/_ref/libra_refmap.so not built (no /root/reference here)")


@pytest.fixture(scope="module")
def mg():
    g = np.load(os.path.join(GOLDEN, "map_cases.npz"))
    return {k: g[k] for k in g.files}


def decode(mg):
    from workload import synth
    return ((mg["raw"].astype(np.float32) + np.float32(synth.DAC_OFFSET)) * (np.float32(synth.DAC_RANGE) / np.float32(synth.DAC_DIGITISATION))).astype(np.float32)


@pytest.fixture(scope="module")
def indexes(mg):
    return {p: build_flat_index([mg["seq0"], mg["seq1"]], mg["pore_vals"], p) for p in ("sensitive", "viral", "fast")}


def test_index_tables_vs_reference_golden(mg, indexes):
    """ora_seq_to_sig + ora_sketch + grouping == ri_idx_gen + ri_idx_dump (rsig.cpp:7-41, rsketch.c:223-274, rawindex.cpp:127-247)."""
    for p, ix in indexes.items():
        assert np.array_equal(table_checksum(ix), mg[f"keys_{p}"]), p
    ix = indexes["sensitive"]
    assert np.array_equal(ix["keys"], mg["keys_full"]) and np.array_equal(ix["key_off"], mg["key_off_full"]) and np.array_equal(ix["vals"], mg["vals_full"])
    assert np.array_equal(bits(ix["fwd"][1]), bits(mg["fwd1"])) and np.array_equal(bits(ix["rev"][1]), bits(mg["rev1"]))


def test_sketch_vs_reference_golden(mg, indexes):
    h, p = MapOracle(indexes["sensitive"]).sketch(mg["sk_ev"])
    assert np.array_equal(h, mg["sk_x"] >> np.uint64(6))                                   # RI_HASH_SHIFT
    assert np.array_equal(p.astype(np.uint64), (mg["sk_y"] & np.uint64(0xFFFFFFFF)) >> np.uint64(1))
    assert np.all((mg["sk_x"] & np.uint64(63)) == np.uint64(6 + 6 - 1))                    # span = k + e - 1


@pytest.mark.parametrize("ci", range(len(CONFIGS)))
def test_map_reads_paf_vs_reference_golden(mg, indexes, ci):
    """Every PAF field and tag (minus mt:f:, the per-read wall time) of every read equals what the reference's
    map_worker_for printed -- mapped/unmapped decision, positions, chunks used, chain counts, scores, MAPQ."""
    preset, kw = CONFIGS[ci]
    mo = MapOracle(indexes[preset])
    res = mo.map_reads(mo.mapopt(**kw), decode(mg), mg["raw_off"], mg["raw_len"])
    want = str(mg["paf"][ci]).split("\n")
    got = mo.paf_lines(res)
    assert len(got) == len(want)
    for r, (a, b) in enumerate(zip(got, want)):
        assert a == b, (r, a, b)
    assert res["mapped"].sum() >= 28 and (res["ci"] > 1).sum() >= 2


def test_chunk_chains_vs_reference_golden(mg, indexes):
    """reg->chains after each of the first chunks of five reads == gen_chains' (through ri_map_frag): order, scores, spans,
    MAPQ, anchors."""
    mo = MapOracle(indexes["sensitive"])
    opt = mo.mapopt(**CONFIGS[0][1])
    sig = decode(mg)
    want_ch = np.ascontiguousarray(mg["chunk_chains"]).view(CHAIN_OUT_DTYPE).reshape(-1)
    want_an = np.ascontiguousarray(mg["chunk_anchors"]).view(ANCHOR_DTYPE).reshape(-1)
    k = 0; co = 0; ao = 0
    for r in (0, 3, 9, 25, 32):
        a = int(mg["raw_off"][r])
        per_chunk, _ = mo.read_chains(opt, sig[a:a + int(mg["raw_len"][r])], max_chunks=6, stop_when_mapped=False)
        for pc in per_chunk:
            n = int(mg["chunk_nchains"][k]); k += 1
            assert len(pc["chains"]) == n
            w = want_ch[co:co + n]; co += n
            for f in ("reference_sequence_index", "start_position", "end_position", "n_anchors", "mapq", "strand"):
                assert np.array_equal(pc["chains"][f], w[f]), f
            assert np.array_equal(bits(pc["chains"]["chaining_score"]), bits(w["chaining_score"]))
            assert np.array_equal(bits(pc["chains"]["alignment_score"]), bits(w["alignment_score"]))
            na = int(w["n_anchors"].sum())
            assert np.array_equal(pc["anchors"], want_an[ao:ao + na]); ao += na
    assert k == len(mg["chunk_nchains"]) and k >= 12


@needs_refmap
def test_map_loop_live_differential(tmp_path):
    """Fresh random data through the compiled reference and through the restatement: identical PAF lines."""
    from oracle import RI_M_DTW_EVALUATE_CHAINS, parse_index_file
    from workload import synth
    rm = RefMap()
    rng = np.random.default_rng(123)
    seqs = [synth.random_reference(150_000, seed=31), synth.random_reference(30_000, seed=32), synth.random_reference(70_000, seed=33)]
    d = str(tmp_path)
    synth.write_fasta(d + "/ref.fa", seqs); synth.write_pore_model(d + "/model.txt")
    rd = synth.simulate_reads(seqs[0], 24, seed=5, max_chunks=30)
    sig = rd.sig.copy()
    for r in range(0, 24, 6):
        a = int(rd.read_off[r]); n = min(int(rd.read_len[r]), 40_000)
        sig[a:a + n] = np.clip(sig[a:a + n] + rng.normal(0, 6.0, n), 31, 199).astype(np.float32)
    for preset, kw in [("sensitive", dict(flag=RI_M_DTW_EVALUATE_CHAINS)), ("fast", dict(flag=RI_M_DTW_EVALUATE_CHAINS, dtw_border_constraint=0, dtw_fill_method=1)),
                       ("viral", dict(flag=0, max_num_chunk=5))]:
        idx = rm.index_build(d + "/ref.fa", d + "/model.txt", preset, dump=d + "/ref.ind")
        ix = parse_index_file(d + "/ref.ind")
        flat = build_flat_index(seqs, rm.pore_model(d + "/model.txt"), preset)
        assert np.array_equal(ix["keys"], flat["keys"]) and np.array_equal(ix["vals"], flat["vals"])
        mo = MapOracle(ix)
        opt = rm.mapopt(**kw)
        paf, mapped, chunks = rm.map_reads(idx, opt, sig, rd.read_off[:-1], rd.read_len, n_threads=4)
        got = mo.paf_lines(mo.map_reads(opt, sig, rd.read_off[:-1], rd.read_len))
        for a, b in zip(got, paf):
            assert a == strip_mt(b)
        rm.index_free(idx)


def test_mapbench_dataset_index_matches_the_restatement(tmp_path):
    """workload/mapbench.py (bench.py's dataset builder): the index file it leaves -- written by the reference's own indexer where
    oracle/_ref/rawalign exists, else from the restatement -- parses to the table oracle.build_flat_index gives for the same
    sequences, and the simulated reads are laid out as the C ABI expects."""
    from oracle import parse_index_file
    from workload import synth
    from workload.mapbench import CONFIGS, Dataset
    ds = Dataset("C1", workdir=str(tmp_path))
    ix = parse_index_file(ds.index_path)
    flat = build_flat_index(ds.seqs, synth.pore_model_as_loaded(), CONFIGS["C1"]["preset"])
    assert np.array_equal(ix["keys"], flat["keys"]) and np.array_equal(ix["key_off"], flat["key_off"]) and np.array_equal(ix["vals"], flat["vals"])
    raw, off, ln, dac_off, dac_scale = ds.reads(16, seed=3)
    assert raw.dtype == np.int16 and len(off) == len(ln) == 16 and int((off.astype(np.int64) + ln).max()) <= len(raw)
    sig = ds.decode(raw)
    assert sig[int(off[0]):int(off[0]) + int(ln[0])].min() > 30.0 and sig[int(off[0]):int(off[0]) + int(ln[0])].max() < 200.0
