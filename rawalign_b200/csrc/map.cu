// map.cu -- the map loop around the hot path, on the device (include/rawalign_b200.h, "ra_map_batch").
//
// Batched form of the reference's per-read map loop (src/rmap.cpp:667-822, map_worker_for): for every chunk round of a
// batch of reads: event detection (K1, through ra_detect_events_batch_dev), seeding (ri_sketch_reg, src/rsketch.c:223-274),
// index lookup (ri_idx_get, src/rawindex.cpp:256-273), anchor sort + chaining DP + traceback (gen_chains,
// src/rmap.cpp:315-507), DTW evaluation of the candidate chains (the hot path, through ra_align_chains_batch_dev),
// gen_primary_chains / comp_mapq (:65-128), is_mapped_with_high_confidence (:594-665) and, when a read leaves the loop, the
// PAF fields of :696-800.  Nothing but the signal goes up and one ra_paf_t per read comes back.
//
// Layout per round (all device-resident):
//   anchors of one read = one segment of 64-bit keys  ref:15 | strand:1 | target:31 | query:17  (the reference sorts every
//   (strand, ref) list by (target, query), rmap.cpp:396-401, and walks the lists ref-major, strand 0 then 1, :429-430: one
//   ascending sort of the packed key gives both), padded to a power of two for the bitonic network;
//   score / predecessor / used flags parallel to the keys; chain anchors (ra_anchor_t, chain END first as traceback_chains
//   emits them, :130-173) in one of two arenas that alternate between rounds, so that the chains kept for re-injection into
//   the next round (:343-357) stay readable while the next round writes its own.
#include "ra_common.cuh"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <thread>

extern "C" void ra__add_launches(ra_ctx *ctx, uint64_t n);        // api.cu
extern "C" int ra__fail(int code, const char *fmt, ...);            // api.cu: sets ra_last_error()
extern "C" int ra__device(const ra_ctx *ctx);
extern "C" int ra__map_sig_whole(const ra_ctx *ctx);
extern "C" int ra__map_host_gather(const ra_ctx *ctx);
extern "C" void ra__swap_resident(ra_ctx *ctx, void **dev_ptr, uint64_t *n_floats);

namespace {

#define MAP_CUDA(call)                                                                                  \
	do {                                                                                                \
		cudaError_t e__ = (call);                                                                       \
		if (e__ != cudaSuccess) return ra__fail(RA_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
	} while (0)

constexpr int kQueryBits = 17, kTargetBits = 31, kRefShift = kQueryBits + kTargetBits + 1;     // 49
constexpr uint32_t kMaxSeq = 1u << (64 - kRefShift);                                           // 32768 sequences
constexpr uint64_t kPadKey = ~0ull;

struct Buf {                                                       // grow-only device buffer
	void *p = nullptr; size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap) return RA_OK;
		if (p) cudaFree(p);
		p = nullptr; cap = 0;
		const size_t want = bytes + bytes / 4 + 256;
		if (cudaMalloc(&p, want) != cudaSuccess) { (void)cudaGetLastError(); p = nullptr; return ra__fail(RA_ENOMEM, "cudaMalloc(%zu) failed", want); }
		cap = want;
		return RA_OK;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct PinBuf {                                                    // grow-only page-locked host buffer; reserve() keeps the first `keep` bytes
	void *p = nullptr; size_t cap = 0;
	int reserve(size_t bytes, size_t keep = 0)
	{
		if (bytes <= cap) return RA_OK;
		size_t want = std::max(bytes, cap + cap / 2);
		void *q = nullptr;
		if (cudaMallocHost(&q, want) != cudaSuccess) { (void)cudaGetLastError(); return ra__fail(RA_ECUDA, "cudaMallocHost(%zu) failed", want); }
		if (p && keep) memcpy(q, p, keep);
		if (p) cudaFreeHost(p);
		p = q; cap = want;
		return RA_OK;
	}
	void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct Slot { uint32_t key, cnt; uint64_t off; };                  // cnt == 0: empty

struct ChainRec {                                                  // ri_chain_t (rmap.h:30-46) with the anchors as an arena offset
	float chaining_score, alignment_score;
	uint32_t ref, start_position, end_position, n_anchors;
	uint32_t mapq, strand;
	uint64_t anchor_off;
};

struct ReadState {
	uint64_t sig_off; uint32_t l_sig;                              // ri_sig_t::sig / l_sig
	uint64_t ev_base; uint32_t n_events;                           // p->events[read_id]
	uint32_t offset;                                               // reg->offset
	uint32_t chunk_start, current_chunk;                           // the loop variables of rmap.cpp:685
	uint32_t n_chains;                                             // reg->n_chains
	uint32_t active, done;
	uint32_t arena;                                                // which chain-anchor arena holds reg->chains' anchors
};

} // namespace

struct ra_index {
	ra_ctx *ctx = nullptr;
	int32_t w = 0, e = 0, n = 0, q = 0, lq = 0, k = 0;
	uint32_t n_seq = 0;
	uint64_t n_keys = 0, n_vals = 0;
	std::vector<uint32_t> seq_len;
	std::vector<std::string> names;
	Buf slots, vals, seq_tab;                                      // seq_tab: per sequence {fwd_off u64, rev_off u64, len u64}
	uint32_t slot_bits = 0;
	// per-context scratch of ra_map_batch lives with the index (one index per context)
	Buf raw16, rd_off, rd_len, rd_offset, rd_scale, sig, state, chains, chains_tmp, paf;
	Buf act, coff, clen, ev_round, evoff_round, ev, seeds_h, seeds_p, seed_cnt, seed_off, n_seed, n_hits, seg_n, seg_base, counters;
	Buf keys, score, pred, used, anc[2], cand, n_cand, cand_base, dtw_chains, group_first, dscore, dkeep, scan_tmp;
	Buf lsig_keep, nt, n_nt, ntf, cxf, bin_base, stage, fin_r, fin_aoff, fin_chain, fin_anc;
	// --dtw-output-cigar results of the last ra_map_batch, host side: per read the range of its alignment in cig_i/j/d
	struct CigDev { Buf parts, first, items, slot_sz, dir_sz, slot_off, dir_off, dirs, si, sj, sd, plen, poff, oi, oj, od, cost, alns, coff, clen, add, ring, cnt; } cg;
	PinBuf cig_i, cig_j, cig_d;                                    // path elements (u32 i, u32 j, f32 diff) of every finished read, in finishing order
	uint64_t cig_total = 0;
	std::vector<uint64_t> cig_start, cig_len, cig_add_i, cig_add_j;   // per read: its range in cig_*; the addend of its last element (global border)
	int n_sm = 0;
	uint32_t tshift = 0, n_bins = 0;                               // bin function of the bucket sort: bin_base[ref << 1 | strand] + (target >> tshift)
	std::vector<uint64_t> last_off; std::vector<uint32_t> last_len; bool last_raw = false;   // the batch left on the device by the last ra_map_batch
	void *pin = nullptr; size_t pin_cap = 0;
	void *pin_sig = nullptr; size_t pin_sig_cap = 0;               // staging of the per-round signal upload from pageable caller memory
	bool sort_attr = false;                                        // the sort kernel's > 48 KB shared memory opt-in (one index = one device)
	cudaStream_t stream = nullptr;
	cudaStream_t stream_hi = nullptr;                              // highest priority: the PCIe gather of a round must not queue behind other batches' kernels
	cudaEvent_t ev_t[16] = {};
};

namespace {

// ---------------------------------------------------------------------------------------------------------------
// host: index file -> flat table -> device open-addressing table
// ---------------------------------------------------------------------------------------------------------------
// sequence table, bin table of the anchor sort, stream and events: everything of an index that follows from the sequence lengths
int finish_index_geometry(ra_ctx *ctx, ra_index *ix, const std::vector<uint64_t> &arena_off)
{
	int rc;
	std::vector<uint64_t> tab(3 * (size_t)ix->n_seq);
	for (uint32_t i = 0; i < ix->n_seq; ++i) { tab[3 * i] = arena_off[2 * i]; tab[3 * i + 1] = arena_off[2 * i + 1]; tab[3 * i + 2] = ix->seq_len[i]; }
	if ((rc = ix->seq_tab.reserve(tab.size() * 8))) return rc;
	MAP_CUDA(cudaMemcpy(ix->seq_tab.p, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice));
	if (2ull * ix->n_seq <= 4096) {                                // bins of the anchor bucket sort (map_kernels.cuh, kBins): uniform over the concatenated reference
		uint32_t ts = 0;
		for (;; ++ts) {
			uint64_t tot = 0;
			for (uint32_t i = 0; i < ix->n_seq; ++i) tot += 2ull * (((uint64_t)ix->seq_len[i] >> ts) + 1);
			if (tot <= 4096) break;
		}
		std::vector<uint32_t> bb(2 * (size_t)ix->n_seq);
		uint32_t acc = 0;
		for (uint32_t i = 0; i < ix->n_seq; ++i)
			for (int st = 0; st < 2; ++st) { bb[2 * i + st] = acc; acc += (uint32_t)(((uint64_t)ix->seq_len[i] >> ts) + 1); }
		if ((rc = ix->bin_base.reserve(bb.size() * 4))) return rc;
		MAP_CUDA(cudaMemcpy(ix->bin_base.p, bb.data(), bb.size() * 4, cudaMemcpyHostToDevice));
		ix->tshift = ts; ix->n_bins = acc;
	}
	MAP_CUDA(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
	{ int lo = 0, hi = 0; MAP_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi)); MAP_CUDA(cudaStreamCreateWithPriority(&ix->stream_hi, cudaStreamNonBlocking, hi)); }
	for (auto &e : ix->ev_t) MAP_CUDA(cudaEventCreate(&e));
	return RA_OK;
}

// arena layout of ra_ref_load: [fwd(0) | rev(0) | fwd(1) | ...], each seq.len floats
std::vector<uint64_t> arena_layout(const ra_index *ix, uint64_t *total)
{
	std::vector<uint64_t> arena_off(2 * (size_t)ix->n_seq);
	uint64_t o = 0;
	for (uint32_t i = 0; i < ix->n_seq; ++i) { arena_off[2 * i] = o; o += ix->seq_len[i]; arena_off[2 * i + 1] = o; o += ix->seq_len[i]; }
	if (total) *total = o;
	return arena_off;
}

// table insertion on the device: one thread per key claims a slot by compare-and-swap on the (key | count << 32) word, linear probing;
// the insertion order (and so the slot a key ends up in) is irrelevant to lookups.  A key seen twice raises *dup.
__global__ void __launch_bounds__(256)
k_ix_insert(const uint32_t *__restrict__ key, const uint32_t *__restrict__ cnt, const uint64_t *__restrict__ off, uint64_t n, Slot *__restrict__ slots,
            uint32_t slot_bits, uint32_t *__restrict__ dup)
{
	const uint32_t m = (1u << slot_bits) - 1u;
	unsigned long long *w = reinterpret_cast<unsigned long long *>(slots);   // 16-byte slots: words 2h (key, cnt) and 2h + 1 (off)
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t k = key[i];
		const unsigned long long word = (unsigned long long)k | ((unsigned long long)cnt[i] << 32);
		uint32_t h = (uint32_t)(k * 0x9E3779B1u) >> (32 - slot_bits);
		for (;;) {
			const unsigned long long prev = atomicCAS(&w[2ull * h], 0ull, word);
			if (prev == 0ull) { w[2ull * h + 1] = off[i]; break; }
			if ((uint32_t)prev == k) { atomicExch(dup, 1u); break; }
			h = (h + 1u) & m;
		}
	}
}

// fwd == NULL: replica (tables of the right size, contents by broadcast).  key / cnt / off: one entry per distinct seed hash,
// off into vals[]; any order.
int build_device_index(ra_ctx *ctx, ra_index *ix, const float *const *fwd, const float *const *rev, const uint32_t *key, const uint32_t *cnt,
                       const uint64_t *off, const uint64_t *vals)
{
	if (ix->n_seq == 0 || ix->n_seq > kMaxSeq) return ra__fail(RA_EUNSUP, "%u reference sequences (supported: 1..%u)", ix->n_seq, kMaxSeq);
	uint64_t total = 0;
	std::vector<uint64_t> arena_off = arena_layout(ix, &total);
	int rc;
	if (fwd) { if ((rc = ra_ref_load(ctx, ix->n_seq, fwd, rev, ix->seq_len.data(), arena_off.data()))) return rc; }
	else if ((rc = ra_ref_alloc(ctx, total))) return rc;
	uint32_t bits = 4;
	while ((1ull << bits) < 2 * ix->n_keys + 16) ++bits;
	if (bits > 31) return ra__fail(RA_EUNSUP, "index has too many distinct seeds (%llu)", (unsigned long long)ix->n_keys);
	ix->slot_bits = bits;
	const uint64_t n_slots = 1ull << bits;
	if ((rc = ix->slots.reserve(n_slots * sizeof(Slot)))) return rc;
	if ((rc = ix->vals.reserve(std::max<uint64_t>(ix->n_vals, 1) * 8))) return rc;
	if (fwd) {
		MAP_CUDA(cudaMemset(ix->slots.p, 0, n_slots * sizeof(Slot)));
		if (ix->n_vals) MAP_CUDA(cudaMemcpy(ix->vals.p, vals, ix->n_vals * 8, cudaMemcpyHostToDevice));
		if (ix->n_keys) {
			Buf dk, dc, dof, dd;
			auto done = [&](int code) { dk.release(); dc.release(); dof.release(); dd.release(); return code; };
			if ((rc = dk.reserve(ix->n_keys * 4)) || (rc = dc.reserve(ix->n_keys * 4)) || (rc = dof.reserve(ix->n_keys * 8)) || (rc = dd.reserve(64))) return done(rc);
			cudaError_t ce;
			if ((ce = cudaMemcpy(dk.p, key, ix->n_keys * 4, cudaMemcpyHostToDevice)) || (ce = cudaMemcpy(dc.p, cnt, ix->n_keys * 4, cudaMemcpyHostToDevice)) ||
			    (ce = cudaMemcpy(dof.p, off, ix->n_keys * 8, cudaMemcpyHostToDevice)) || (ce = cudaMemset(dd.p, 0, 4)))
				return done(ra__fail(RA_ECUDA, "index upload: %s", cudaGetErrorString(ce)));
			k_ix_insert<<<(unsigned)std::min<uint64_t>(148 * 16, (ix->n_keys + 255) / 256), 256>>>(dk.as<uint32_t>(), dc.as<uint32_t>(), dof.as<uint64_t>(), ix->n_keys,
			                                                                                    ix->slots.as<Slot>(), bits, dd.as<uint32_t>());
			uint32_t dup = 0;
			if ((ce = cudaMemcpy(&dup, dd.p, 4, cudaMemcpyDeviceToHost)) || (ce = cudaGetLastError())) return done(ra__fail(RA_ECUDA, "index insert: %s", cudaGetErrorString(ce)));
			done(0);
			if (dup) return ra__fail(RA_EINVAL, "index table holds a seed hash twice");
			ra__add_launches(ctx, 1);
		}
	}
	return finish_index_geometry(ctx, ix, arena_off);
}

} // namespace

extern "C" {

void ra_ri_mapopt_init(ra_ri_mapopt_t *o)
{
	memset(o, 0, sizeof(*o));
	o->bp_per_sec = 450; o->sample_rate = 4000; o->chunk_size = 4000;
	o->max_gap_length = 2000; o->max_target_gap_length = 5000; o->chaining_band_length = 5000; o->max_num_skips = 25;
	o->min_num_anchors = 2; o->num_best_chains = 3; o->min_chaining_score = 10.0f;
	o->step_size = 1; o->min_events = 50; o->max_num_chunk = 30; o->min_chain_anchor = 2; o->min_chain_anchor_out = 2;
	o->min_bestmap_ratio = 1.2f; o->min_bestmap_ratio_out = 1.2f; o->min_meanmap_ratio = 5; o->min_meanmap_ratio_out = 5;
	o->mini_batch_size = 500000000;
	o->window_length1 = 3; o->window_length2 = 6; o->threshold1 = 4.30265f; o->threshold2 = 2.57058f; o->peak_height = 1.0f;
	o->t_threshold = 1.5f; o->tn_samples = 5; o->ttest_freq = 500; o->tmin_reads = 500;
	o->dtw_border_constraint = RA_DTW_BORDER_CONSTRAINT_SPARSE; o->dtw_fill_method = RA_DTW_FILL_METHOD_BANDED;
	o->dtw_band_radius_frac = 0.10f; o->dtw_match_bonus = 0.4f; o->dtw_min_score = 20.0f;
}

int ra_index_load_flat(ra_ctx *ctx, int32_t w, int32_t e, int32_t n, int32_t q, int32_t lq, int32_t k, uint32_t n_seq, const uint32_t *seq_len,
                       const float *const *fwd, const float *const *rev, uint64_t n_keys, const uint64_t *keys, const uint64_t *key_off,
                       const uint64_t *vals, ra_index **out)
{
	if (!ctx || !out || !seq_len || !fwd || !rev || (n_keys && (!keys || !key_off || !vals))) return ra__fail(RA_EINVAL, "NULL argument");
	if (e < 2 || e > 9 || (lq + 2) * e > 63 || q < 1 || q > 32 || lq < 0 || lq > q) return ra__fail(RA_EINVAL, "bad sketch parameters e=%d q=%d lq=%d", e, q, lq);
	ra_index *ix = new ra_index();
	ix->ctx = ctx; ix->w = w; ix->e = e; ix->n = n; ix->q = q; ix->lq = lq; ix->k = k; ix->n_seq = n_seq; ix->n_keys = n_keys;
	ix->n_vals = n_keys ? key_off[n_keys] : 0;
	ix->seq_len.assign(seq_len, seq_len + n_seq);
	std::vector<uint32_t> k32(n_keys), c32(n_keys);
	for (uint64_t i = 0; i < n_keys; ++i) {
		const uint64_t c = key_off[i + 1] - key_off[i];
		if (keys[i] >> 32 || c == 0 || c >> 32 || key_off[i + 1] > ix->n_vals) {
			const uint64_t bad_key = keys[i];
			ra_index_destroy(ix);
			return ra__fail(RA_EINVAL, "index table entry %llu: hash %llu must fit 32 bits (rsketch.c masks hash64 to 32 bits) and own 1..2^32-1 in-range positions",
			                (unsigned long long)i, (unsigned long long)bad_key);
		}
		k32[i] = (uint32_t)keys[i]; c32[i] = (uint32_t)c;
	}
	int rc = build_device_index(ctx, ix, fwd, rev, k32.data(), c32.data(), key_off, vals);
	if (rc) { ra_index_destroy(ix); return rc; }
	*out = ix;
	return RA_OK;
}

// rawindex.cpp:275-315 (ri_idx_dump) read back: see the header comment.
int ra_index_load_file(ra_ctx *ctx, const char *path, ra_index **out)
{
	if (!ctx || !path || !out) return ra__fail(RA_EINVAL, "NULL argument");
	FILE *f = fopen(path, "rb");
	if (!f) return ra__fail(RA_EINVAL, "cannot open index file %s", path);
	auto bad = [&](const char *what) { fclose(f); return ra__fail(RA_EINVAL, "index file %s: %s", path, what); };
	char magic[2];
	uint32_t pars[8];
	if (fread(magic, 1, 2, f) != 2 || magic[0] != 'R' || magic[1] != 'I') return bad("bad magic");
	if (fread(pars, 4, 8, f) != 8) return bad("truncated header");
	const uint32_t n_seq = pars[6];
	std::vector<uint32_t> seq_len(n_seq);
	std::vector<std::vector<float>> fw(n_seq), rv(n_seq);
	std::vector<std::string> names(n_seq);
	for (uint32_t i = 0; i < n_seq; ++i) {
		uint8_t l;
		if (fread(&l, 1, 1, f) != 1) return bad("truncated sequence table");
		names[i].resize(l);
		if (l && fread(&names[i][0], 1, l, f) != l) return bad("truncated sequence name");
		if (fread(&seq_len[i], 4, 1, f) != 1) return bad("truncated sequence length");
		fw[i].resize(seq_len[i]); rv[i].resize(seq_len[i]);
		if (fread(fw[i].data(), 4, seq_len[i], f) != seq_len[i] || fread(rv[i].data(), 4, seq_len[i], f) != seq_len[i]) return bad("truncated signal");
	}
	const int b = 14;                                              // ri_idx_load: ri_idx_init(14, ...), rawindex.cpp:330
	std::vector<uint32_t> k32, c32;                               // one entry per distinct seed hash, in file order
	std::vector<uint64_t> koff;
	std::vector<uint64_t> pool;                                    // every bucket's p[] and the single positions, in file order
	for (uint32_t i = 0; i < (1u << b); ++i) {
		int32_t nb; uint32_t size;
		if (fread(&nb, 4, 1, f) != 1 || nb < 0) return bad("truncated bucket");
		const uint64_t pbase = pool.size();
		pool.resize(pbase + (size_t)nb);
		if (nb && fread(pool.data() + pbase, 8, (size_t)nb, f) != (size_t)nb) return bad("truncated bucket positions");
		if (fread(&size, 4, 1, f) != 1) return bad("truncated bucket");
		for (uint32_t j = 0; j < size; ++j) {
			uint64_t x[2];
			if (fread(x, 8, 2, f) != 2) return bad("truncated hash table");
			const uint64_t hash = ((x[0] >> 1) << b) | i;          // key = hash >> b << 1 | single (rawindex.cpp:222-226)
			if (hash >> 32) return bad("seed hash does not fit 32 bits (rsketch.c masks hash64 to 32 bits)");
			k32.push_back((uint32_t)hash);
			if (x[0] & 1) { pool.push_back(x[1]); koff.push_back(pool.size() - 1); c32.push_back(1); }
			else {
				const uint64_t st = x[1] >> 32, c = (uint32_t)x[1];
				if (c == 0 || st + c > (uint64_t)nb) return bad("hash table entry points outside its bucket");
				koff.push_back(pbase + st); c32.push_back((uint32_t)c);
			}
		}
	}
	fclose(f);
	const uint32_t e = pars[1], q = pars[3], lq = pars[4];
	if (e < 2 || e > 9 || (lq + 2) * e > 63 || q < 1 || q > 32 || lq > q) return ra__fail(RA_EINVAL, "index file %s: bad sketch parameters e=%u q=%u lq=%u", path, e, q, lq);
	std::vector<const float *> fp(n_seq), rp(n_seq);
	for (uint32_t i = 0; i < n_seq; ++i) { fp[i] = fw[i].data(); rp[i] = rv[i].data(); }
	ra_index *ix = new ra_index();
	ix->ctx = ctx; ix->w = (int32_t)pars[0]; ix->e = (int32_t)e; ix->n = (int32_t)pars[2]; ix->q = (int32_t)q; ix->lq = (int32_t)lq; ix->k = (int32_t)pars[5];
	ix->n_seq = n_seq; ix->n_keys = k32.size(); ix->n_vals = pool.size(); ix->seq_len = seq_len; ix->names = names;
	// the table is keyed by hash and the positions of a key are contiguous in pool[] (ascending, rawindex.cpp:232): no sorting needed
	int rc = build_device_index(ctx, ix, fp.data(), rp.data(), k32.data(), c32.data(), koff.data(), pool.data());
	if (rc) { ra_index_destroy(ix); return rc; }
	*out = ix;
	return RA_OK;
}

void ra_index_destroy(ra_index *ix)
{
	if (!ix) return;
	Buf *all[] = { &ix->slots, &ix->vals, &ix->seq_tab, &ix->raw16, &ix->rd_off, &ix->rd_len, &ix->rd_offset, &ix->rd_scale, &ix->sig, &ix->state, &ix->chains,
	               &ix->chains_tmp, &ix->paf, &ix->act, &ix->coff, &ix->clen, &ix->ev_round, &ix->evoff_round, &ix->ev, &ix->seeds_h, &ix->seeds_p,
	               &ix->seed_cnt, &ix->seed_off, &ix->n_seed, &ix->n_hits, &ix->seg_n, &ix->seg_base, &ix->counters, &ix->keys, &ix->score, &ix->pred,
	               &ix->used, &ix->anc[0], &ix->anc[1], &ix->cand, &ix->n_cand, &ix->cand_base, &ix->dtw_chains, &ix->group_first, &ix->dscore,
	               &ix->dkeep, &ix->scan_tmp, &ix->lsig_keep, &ix->nt, &ix->n_nt, &ix->ntf, &ix->cxf, &ix->bin_base, &ix->stage, &ix->fin_r, &ix->fin_aoff, &ix->fin_chain, &ix->fin_anc };
	for (Buf *b : all) b->release();
	if (ix->pin) cudaFreeHost(ix->pin);
	if (ix->pin_sig) cudaFreeHost(ix->pin_sig);
	ix->cig_i.release(); ix->cig_j.release(); ix->cig_d.release();
	{ ra_index::CigDev &G = ix->cg; Buf *cg[] = { &G.parts, &G.first, &G.items, &G.slot_sz, &G.dir_sz, &G.slot_off, &G.dir_off, &G.dirs, &G.si, &G.sj, &G.sd, &G.plen, &G.poff, &G.oi, &G.oj, &G.od, &G.cost, &G.alns, &G.coff, &G.clen, &G.add, &G.ring, &G.cnt }; for (Buf *b : cg) b->release(); }
	if (ix->stream) cudaStreamDestroy(ix->stream);
	if (ix->stream_hi) cudaStreamDestroy(ix->stream_hi);
	for (auto &e : ix->ev_t) if (e) cudaEventDestroy(e);
	delete ix;
}

/* ---- replication across GPUs: rank 0 loads the index, the others allocate one of the same geometry and receive the tables --- */
int ra_index_tables(ra_index *ix, void **ptrs, uint64_t *bytes)
{
	if (!ix || !ptrs || !bytes) return ra__fail(RA_EINVAL, "NULL argument");
	float *ref = nullptr; uint64_t ref_n = 0;
	int rc = ra_ref_device(ix->ctx, &ref, &ref_n);
	if (rc) return rc;
	ptrs[0] = ix->slots.p; bytes[0] = (1ull << ix->slot_bits) * sizeof(Slot);
	ptrs[1] = ix->vals.p;  bytes[1] = ix->n_vals * 8;
	ptrs[2] = ref;         bytes[2] = ref_n * 4;
	return RA_OK;
}

int ra_index_alloc(ra_ctx *ctx, const int32_t *info /* the 9 values of ra_index_info */, uint64_t n_keys, uint64_t n_vals, const uint32_t *seq_len, ra_index **out)
{
	if (!ctx || !info || !seq_len || !out) return ra__fail(RA_EINVAL, "NULL argument");
	ra_index *ix = new ra_index();
	ix->ctx = ctx; ix->w = info[0]; ix->e = info[1]; ix->n = info[2]; ix->q = info[3]; ix->lq = info[4]; ix->k = info[5]; ix->n_seq = (uint32_t)info[6];
	ix->n_keys = n_keys; ix->n_vals = n_vals;
	ix->seq_len.assign(seq_len, seq_len + ix->n_seq);
	int rc = build_device_index(ctx, ix, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);     // geometry and tables of the right size, contents by broadcast
	if (rc) { ra_index_destroy(ix); return rc; }
	*out = ix;
	return RA_OK;
}

// A second index with the same contents for another context (another batch in flight on the same GPU, or a peer GPU of the same
// process): tables copied device to device, no host parse.
int ra_index_clone(ra_index *src, ra_ctx *ctx, ra_index **out)
{
	if (!src || !ctx || !out) return ra__fail(RA_EINVAL, "NULL argument");
	ra_index *ix = new ra_index();
	ix->ctx = ctx; ix->w = src->w; ix->e = src->e; ix->n = src->n; ix->q = src->q; ix->lq = src->lq; ix->k = src->k; ix->n_seq = src->n_seq;
	ix->seq_len = src->seq_len; ix->names = src->names; ix->n_vals = src->n_vals;
	ix->n_keys = ((1ull << src->slot_bits) - 16) / 2;               // the key count that sizes the same table; the true count is set below
	MAP_CUDA(cudaSetDevice(ra__device(ctx)));
	int rc = build_device_index(ctx, ix, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
	if (rc) { ra_index_destroy(ix); return rc; }
	ix->n_keys = src->n_keys;
	void *sp[3], *dp[3]; uint64_t sb[3], db[3];
	if ((rc = ra_index_tables(src, sp, sb)) || (rc = ra_index_tables(ix, dp, db))) { ra_index_destroy(ix); return rc; }
	for (int i = 0; i < 3; ++i) {
		if (sb[i] != db[i]) { ra_index_destroy(ix); return ra__fail(RA_EINVAL, "clone geometry differs (table %d: %llu vs %llu bytes)", i, (unsigned long long)sb[i], (unsigned long long)db[i]); }
		if (sb[i]) {
			cudaError_t ce = cudaMemcpyPeer(dp[i], ra__device(ctx), sp[i], ra__device(src->ctx), sb[i]);
			if (ce != cudaSuccess) { ra_index_destroy(ix); return ra__fail(RA_ECUDA, "index clone: %s", cudaGetErrorString(ce)); }
		}
	}
	MAP_CUDA(cudaDeviceSynchronize());
	*out = ix;
	return RA_OK;
}

int ra_index_info(const ra_index *ix, int32_t *out, uint32_t *seq_len_out)
{
	if (!ix || !out) return ra__fail(RA_EINVAL, "NULL argument");
	out[0] = ix->w; out[1] = ix->e; out[2] = ix->n; out[3] = ix->q; out[4] = ix->lq; out[5] = ix->k; out[6] = (int32_t)ix->n_seq;
	out[7] = (int32_t)ix->n_keys; out[8] = (int32_t)ix->n_vals;
	if (seq_len_out) memcpy(seq_len_out, ix->seq_len.data(), ix->n_seq * 4);
	return RA_OK;
}

} // extern "C"

#include "map_kernels.cuh"
#include "map_chain.cuh"
#include "map_cigar.cuh"
#include "map_batch.cuh"
#include "map_index_build.cuh"
#include "map_index_host.cuh"
