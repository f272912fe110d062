// map_kernels.cuh -- device kernels of the map loop (included by map.cu only).
#pragma once

namespace mapk {

struct Params {                                                    // what the kernels read from ri_mapopt_t / ri_idx_t
	uint32_t chunk_size, max_num_chunk, min_events;
	int32_t max_gap_length, max_target_gap_length, chaining_band_length, max_num_skips, min_num_anchors, num_best_chains;
	float min_chaining_score;
	uint32_t min_chain_anchor;
	float min_bestmap_ratio, min_meanmap_ratio, dtw_min_score;
	uint32_t sample_rate, bp_per_sec;
	uint32_t dtw, no_filter;                                       // RI_M_DTW_EVALUATE_CHAINS, RI_M_DISABLE_CHAININGSCORE_FILTERING
	int32_t e, q, lq;
	uint32_t n_seq, cap_chains, slot_bits;
};

__device__ __forceinline__ uint64_t hash64(uint64_t key, uint64_t mask)                  // rsketch.c:6-15
{
	key = (~key + (key << 21)) & mask;
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8)) & mask;
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4)) & mask;
	key = key ^ key >> 28;
	key = (key + (key << 31)) & mask;
	return key;
}

__device__ __forceinline__ uint64_t pack_key(uint32_t ref, uint32_t strand, uint32_t target, uint32_t query)
{
	return ((uint64_t)ref << kRefShift) | ((uint64_t)strand << (kQueryBits + kTargetBits)) | ((uint64_t)target << kQueryBits) | (uint64_t)query;
}
__device__ __forceinline__ uint32_t key_target(uint64_t k) { return (uint32_t)(k >> kQueryBits) & 0x7fffffffu; }
__device__ __forceinline__ uint32_t key_query(uint64_t k) { return (uint32_t)k & ((1u << kQueryBits) - 1u); }
__device__ __forceinline__ uint32_t key_list(uint64_t k) { return (uint32_t)(k >> (kQueryBits + kTargetBits)); }   // ref << 1 | strand

// ---------------------------------------------------------------------------------------------------------------
// round set-up: the reads still in the chunk loop of rmap.cpp:685, in read order, and their next chunk
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
k_round_setup(ReadState *__restrict__ st, uint32_t n_reads, Params P, uint32_t *__restrict__ act, uint64_t *__restrict__ coff,
              uint32_t *__restrict__ clen, uint32_t *__restrict__ counters /* [0] n_act, [1] max chunk length */)
{
	__shared__ uint32_t warp_tot[32];
	__shared__ uint32_t base_s, maxlen_s;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	if (tid == 0) { base_s = 0; maxlen_s = 0; }
	__syncthreads();
	for (uint32_t r0 = 0; r0 < n_reads; r0 += 1024) {
		const uint32_t r = r0 + tid;
		const bool on = r < n_reads && st[r].active;
		const uint32_t m = __ballot_sync(0xffffffffu, on);
		if (lane == 0) warp_tot[wid] = __popc(m);
		__syncthreads();
		uint32_t before = 0, total = 0;
		for (int w = 0; w < 32; ++w) { const uint32_t t = warp_tot[w]; if (w < wid) before += t; total += t; }
		const uint32_t base = base_s;
		if (on) {
			const uint32_t i = base + before + __popc(m & ((1u << lane) - 1u));
			const ReadState s = st[r];
			const uint32_t len = min(P.chunk_size, s.l_sig - s.chunk_start);                      // rmap.cpp:686-687
			act[i] = r; coff[i] = s.sig_off + s.chunk_start; clen[i] = len;
			atomicMax(&maxlen_s, len);
			atomicAdd(reinterpret_cast<unsigned long long *>(counters + 14), (unsigned long long)len);   // [14..15]: samples of the round
		}
		__syncthreads();
		if (tid == 0) base_s = base + total;
		__syncthreads();
	}
	if (tid == 0) { counters[0] = base_s; counters[1] = maxlen_s; }
}

// events of the round's chunks (packed CSR over the active list) appended to each read's event array, rmap.cpp:554-567
__global__ void __launch_bounds__(128)
k_append_events(ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, const float *__restrict__ ev_round,
                const uint64_t *__restrict__ evoff, float *__restrict__ ev)
{
	const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (i >= n_act) return;
	const uint32_t r = act[i];
	const uint64_t a = evoff[i];
	const uint32_t n = (uint32_t)(evoff[i + 1] - a);
	float *dst = ev + st[r].ev_base + st[r].n_events;
	for (uint32_t k = lane; k < n; k += 32) dst[k] = ev_round[a + k];
	__syncwarp();
	if (lane == 0) st[r].n_events += n;
}

// ---------------------------------------------------------------------------------------------------------------
// seeding: ri_sketch_reg (rsketch.c:223-274) + ri_idx_get (rawindex.cpp:256-273), one warp per read
// ---------------------------------------------------------------------------------------------------------------
// The accept filter (:243: skip an event closer than LAST_SIG_DIFF to the last ACCEPTED one) is a recurrence; the warp
// resolves it 32 events at a time in registers (every lane replays the same 32 steps on shuffled values), the 5-bit
// quantised symbols of the accepted events go to shared memory, and every accepted event of rank >= e-1 hashes the last e
// symbols (:254-255).  Seeds land at seeds_*[evoff[i] + s] (there are fewer seeds than events).
__global__ void __launch_bounds__(32)
k_seed(const ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, const float *__restrict__ ev_round,
       const uint64_t *__restrict__ evoff, Params P, const Slot *__restrict__ slots, const ChainRec *__restrict__ chains,
       uint32_t *__restrict__ seeds_h, uint32_t *__restrict__ seeds_p, uint32_t *__restrict__ seed_cnt, uint64_t *__restrict__ seed_off,
       uint32_t *__restrict__ n_seed_out, uint32_t *__restrict__ seg_n /* anchors of the read this round */, uint32_t *__restrict__ seg_pad,
       uint32_t *__restrict__ skip, unsigned long long *__restrict__ totals /* [0] seeds, [1] anchors, [2] longest segment */)
{
	extern __shared__ uint8_t sym[];                               // quantised symbols of the accepted events
	const uint32_t i = blockIdx.x;
	const int lane = threadIdx.x;
	if (i >= n_act) return;
	const uint32_t r = act[i];
	const uint64_t a = evoff[i];
	const uint32_t n_ev = (uint32_t)(evoff[i + 1] - a);
	const float *__restrict__ ev = ev_round + a;
	uint32_t prev = 0;
	{
		const uint32_t nc = st[r].n_chains;
		for (uint32_t c = lane; c < nc; c += 32) prev += chains[(size_t)r * P.cap_chains + c].n_anchors;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) prev += __shfl_xor_sync(0xffffffffu, prev, o);
	}
	if (n_ev < P.min_events) {                                     // rmap.cpp:569-572: no chaining this round; reg->chains stay
		if (lane == 0) { n_seed_out[i] = 0; seg_n[i] = prev; seg_pad[i] = prev; skip[i] = 1; }
		return;
	}
	const uint32_t quant_bit = (uint32_t)P.lq + 2u, shift_r = 32u - (uint32_t)P.q;
	const uint32_t mask_l_quant = (1u << P.lq) - 1u;
	const uint64_t mask = (1ull << 32) - 1, mask_events = (1ull << (quant_bit * (uint32_t)P.e)) - 1;
	float last = 0.0f;
	uint32_t n_acc = 0;
	for (uint32_t b = 0; b < n_ev; b += 32) {
		const uint32_t idx = b + lane;
		const float v = idx < n_ev ? ev[idx] : 0.0f;
		const uint32_t cnt = min(32u, n_ev - b);
		uint32_t accmask = 0;
		for (uint32_t t = 0; t < cnt; ++t) {
			const float x = __shfl_sync(0xffffffffu, v, (int)t);
			// fabs(s[i] - s[l_sigpos]) < LAST_SIG_DIFF in double after an fp32 subtraction == fabsf(d) < 0.3f; RI_MASK_SIGNAL :243
			const bool rej = ((b + t) > 0 && fabsf(__fsub_rn(x, last)) < 0.3f) || x == 3.402823466e+32f;
			if (!rej) { last = x; accmask |= 1u << t; }
		}
		const bool mine = (accmask >> lane) & 1u;
		const uint32_t rank = n_acc + __popc(accmask & ((1u << lane) - 1u));
		if (mine) {
			const uint32_t s = __float_as_uint(v);
			sym[rank] = (uint8_t)((s >> 30 << P.lq) | ((s >> shift_r) & mask_l_quant));             // :247
		}
		__syncwarp();
		if (mine && rank + 1u >= (uint32_t)P.e) {
			uint64_t qv = 0;
			for (int j = P.e - 1; j >= 0; --j) qv = (qv << quant_bit) | sym[rank - (uint32_t)j];    // oldest symbol highest, :254
			qv &= mask_events;
			const uint32_t sidx = rank + 1u - (uint32_t)P.e;
			seeds_h[a + sidx] = (uint32_t)hash64(qv, mask);
			seeds_p[a + sidx] = idx;
		}
		n_acc += __popc(accmask);
		__syncwarp();
	}
	const uint32_t n_seed = n_acc + 1u >= (uint32_t)P.e ? n_acc + 1u - (uint32_t)P.e : 0u;
	__syncwarp();
	uint32_t hits = 0;
	const uint32_t nslots_mask = (1u << P.slot_bits) - 1u;
	for (uint32_t s = lane; s < n_seed; s += 32) {
		const uint32_t key = seeds_h[a + s];
		uint32_t h = (uint32_t)(key * 0x9E3779B1u) >> (32 - P.slot_bits);
		uint32_t c = 0; uint64_t off = 0;
		for (;;) {
			const Slot sl = slots[h];
			if (sl.cnt == 0) break;
			if (sl.key == key) { c = sl.cnt; off = sl.off; break; }
			h = (h + 1) & nslots_mask;
		}
		seed_cnt[a + s] = c; seed_off[a + s] = off;
		hits += c;
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) hits += __shfl_xor_sync(0xffffffffu, hits, o);
	if (lane == 0) {
		const uint32_t n = hits + prev;
		uint32_t pad = 1; while (pad < n) pad <<= 1;
		if (n == 0) pad = 0;
		n_seed_out[i] = n_seed; seg_n[i] = n; seg_pad[i] = pad; skip[i] = 0;
		atomicAdd(totals, (unsigned long long)n_seed); atomicAdd(totals + 1, (unsigned long long)n); atomicMax(totals + 2, (unsigned long long)n);
	}
}

// anchors of the round: the previous chunk's chains re-injected (rmap.cpp:343-357) + one per seed hit (:383-390), as packed
// keys in the read's (power-of-two padded) segment.  Reads that skipped chaining carry their chains' anchors into the new
// arena unchanged.
__global__ void __launch_bounds__(32)
k_fill_anchors(ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, const uint64_t *__restrict__ evoff, Params P,
               ChainRec *__restrict__ chains, const ra_anchor_t *__restrict__ anc_prev0, const ra_anchor_t *__restrict__ anc_prev1,
               ra_anchor_t *__restrict__ anc_cur, uint32_t cur_arena, const uint32_t *__restrict__ seeds_p, const uint32_t *__restrict__ seed_cnt,
               const uint64_t *__restrict__ seed_off, const uint64_t *__restrict__ vals, const uint32_t *__restrict__ n_seed_in,
               const uint32_t *__restrict__ seg_n, const uint32_t *__restrict__ seg_pad, const uint64_t *__restrict__ seg_base,
               const uint32_t *__restrict__ skip, uint64_t *__restrict__ keys)
{
	const uint32_t i = blockIdx.x;
	const int lane = threadIdx.x;
	if (i >= n_act) return;
	const uint32_t r = act[i];
	const uint64_t base = seg_base[i];
	const ReadState s = st[r];
	const ra_anchor_t *__restrict__ prev = s.arena ? anc_prev1 : anc_prev0;
	ChainRec *rc = chains + (size_t)r * P.cap_chains;
	if (skip[i]) {
		uint64_t o = base;
		for (uint32_t c = 0; c < s.n_chains; ++c) {
			const ChainRec ch = rc[c];
			for (uint32_t k = lane; k < ch.n_anchors; k += 32) anc_cur[o + k] = prev[ch.anchor_off + k];
			__syncwarp();
			if (lane == 0) rc[c].anchor_off = o;
			o += ch.n_anchors;
		}
		if (lane == 0) st[r].arena = cur_arena;
		return;
	}
	uint64_t o = base;
	for (uint32_t c = 0; c < s.n_chains; ++c) {
		const ChainRec ch = rc[c];
		for (uint32_t k = lane; k < ch.n_anchors; k += 32) {
			const ra_anchor_t an = prev[ch.anchor_off + k];
			keys[o + k] = pack_key(ch.ref, ch.strand, an.target_position, an.query_position);
		}
		o += ch.n_anchors;
	}
	const uint64_t a = evoff[i];
	const uint32_t n_seed = n_seed_in[i];
	const uint32_t chunk_start = s.offset;                                                          // gen_chains(..., reg->offset, ...) :574
	// hits of 32 seeds at a time, spread evenly over the lanes (the hit counts are heavy-tailed: a lane that owned one
	// frequent seed would hold the warp): hit h of the tile belongs to the seed whose exclusive offset is the last one <= h
	__shared__ uint32_t s_excl[32], s_q[32];
	__shared__ uint64_t s_vo[32];
	for (uint32_t b = 0; b < n_seed; b += 32) {
		const uint32_t sd = b + lane;
		const uint32_t c = sd < n_seed ? seed_cnt[a + sd] : 0u;
		uint32_t inc = c;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += y; }
		const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
		s_excl[lane] = inc - c;
		s_vo[lane] = sd < n_seed ? seed_off[a + sd] : 0ull;
		s_q[lane] = sd < n_seed ? seeds_p[a + sd] + chunk_start : 0u;                                  // pi + chunk_start, :387
		__syncwarp();
		for (uint32_t h = lane; h < total; h += 32) {
			int lo = 0;
#pragma unroll
			for (int step = 16; step > 0; step >>= 1) if (s_excl[lo + step] <= h) lo += step;      // seeds without hits share an offset: the last one wins
			const uint64_t kv = vals[s_vo[lo] + (h - s_excl[lo])];
			keys[o + h] = pack_key((uint32_t)(kv >> 32), (uint32_t)(kv & 1), (uint32_t)(kv >> 1) & 0x7fffffffu, s_q[lo]);   // :385-387
		}
		__syncwarp();
		o += total;
	}
	for (uint64_t k = base + seg_n[i] + lane; k < base + seg_pad[i]; k += 32) keys[k] = kPadKey;
}

// ---------------------------------------------------------------------------------------------------------------
// anchor sort (rmap.cpp:396-401): one CTA per read, bitonic network over the padded segment
// ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t kSortSmem = 16384;                              // keys staged in shared memory (128 KB, opt-in)
constexpr int kSortThreads = 1024;
__global__ void __launch_bounds__(kSortThreads)
k_sort_segments(uint64_t *__restrict__ keys, const uint64_t *__restrict__ seg_base, const uint32_t *__restrict__ seg_pad,
                const uint32_t *__restrict__ skip, uint32_t n_act, const uint32_t *__restrict__ seg_n, uint32_t done_n /* segments of up to this many anchors were sorted by k_sort_buckets */)
{
	extern __shared__ uint64_t sk[];
	const uint32_t i = blockIdx.x;
	if (i >= n_act || skip[i]) return;
	const uint32_t n = seg_pad[i];
	if (n < 2 || seg_n[i] <= done_n) return;
	uint64_t *g = keys + seg_base[i];
	const int tid = threadIdx.x;
	const bool in_smem = n <= kSortSmem;
	uint64_t *w = in_smem ? sk : g;                                // longer segments are sorted in place in global memory
	if (in_smem) {
		for (uint32_t k = tid; k < n; k += kSortThreads) sk[k] = g[k];
		__syncthreads();
	}
	for (uint32_t size = 2; size <= n; size <<= 1)
		for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
			for (uint32_t t = tid; t < (n >> 1); t += kSortThreads) {
				const uint32_t lo = 2 * t - (t & (stride - 1));
				const uint32_t hi = lo + stride;
				const bool up = (lo & size) == 0;
				const uint64_t x = w[lo], y = w[hi];
				const bool sw = (x > y) == up;
				w[lo] = sw ? y : x; w[hi] = sw ? x : y;
			}
			__syncthreads();
		}
	if (in_smem)
		for (uint32_t k = tid; k < n; k += kSortThreads) g[k] = sk[k];
}

// ---------------------------------------------------------------------------------------------------------------
// anchor sort, bucket form: one CTA per read, everything in shared memory
// ---------------------------------------------------------------------------------------------------------------
// The packed keys are uniformly spread over the reference (seed hits are), so one counting pass over a monotone bin function
// -- bin = bin_base[ref << 1 | strand] + (target >> tshift), at most kBins bins over the whole index -- leaves a handful of
// keys per bin; the bins are then finished independently (insertion sort by one thread, or a warp-wide bitonic pass for the
// rare crowded bin, e.g. the read's true locus).  O(n) work per read instead of the bitonic network's n log^2 n.
constexpr uint32_t kBins = 4096;
constexpr uint32_t kBucketMaxKeys = 8192;                          // the common size class: three CTAs per SM
constexpr uint32_t kBucketMaxKeysL = 20480;                        // long segments (reads with several chunks of anchors): one CTA per SM
constexpr int kBucketThreads = 512;
constexpr uint32_t kGroupKeys = 4096;                              // MAXK == 0: keys of one run of bins finished in shared memory
constexpr uint32_t kFineBins = 2048;
constexpr int kScanIters = (int)(kBins / kBucketThreads);         // bins per thread in the CTA-wide scan
constexpr uint32_t kMaxCrowded = 63;
constexpr uint32_t kSmallBin = 24;                                 // bins up to this size are finished by one thread                               // MAXK == 0: the run is binned again, up to 2^F times finer
template <uint32_t MAXK>
struct BucketSmemT {
	uint64_t keys[MAXK ? MAXK : kGroupKeys];                       // MAXK == 0: the binned keys are staged in a global-memory arena; keys[] holds one run of bins at a time
	uint64_t keys2[MAXK ? 1 : kGroupKeys];                         // MAXK == 0: the run after the second binning
	uint32_t fine[MAXK ? 1 : kFineBins + 1];
	uint32_t count[kBins + 1];
	uint32_t cursor[kBins];
	uint32_t warp_tot[kScanIters * (kBucketThreads / 32)];
	uint32_t crowd[1 + kMaxCrowded];                               // [0]: number of bins with more than kSmallBin keys, then their indices
};

// exclusive scan of a[0 .. n), n <= kBins, in place by the whole CTA (a[n] is not written); cur[] receives a copy.  Thread t holds
// bins t, t + 512, ...: warp-inclusive scans in registers, one pass of warp 0 over the 128 warp totals, three barriers in all.
__device__ __forceinline__ void cta_exclusive_scan(uint32_t *a, uint32_t *cur, uint32_t n, uint32_t *warp_tot, int tid, int lane, int wid,
                                                   uint32_t *crowd = nullptr /* bins of more than kSmallBin keys are appended (crowd[0] zeroed by the caller) */)
{
	constexpr int W = kBucketThreads / 32;
	uint32_t v[kScanIters], inc[kScanIters];
#pragma unroll
	for (int i = 0; i < kScanIters; ++i) {
		const uint32_t b = (uint32_t)(i * kBucketThreads + tid);
		v[i] = b < n ? a[b] : 0u;
		uint32_t x = v[i];
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
		inc[i] = x;
		if (lane == 31) warp_tot[i * W + wid] = x;                 // entry i * W + wid covers bins i * 512 + wid * 32 + [0, 32): bin order
	}
	__syncthreads();
	if (wid == 0) {                                                // exclusive scan of the kScanIters * W totals, (kScanIters * W / 32) per lane
		constexpr int PER = kScanIters * W / 32;
		uint32_t t[PER], sum = 0;
#pragma unroll
		for (int k = 0; k < PER; ++k) { t[k] = warp_tot[PER * lane + k]; sum += t[k]; }
		uint32_t x = sum;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
		uint32_t ex = x - sum;
#pragma unroll
		for (int k = 0; k < PER; ++k) { warp_tot[PER * lane + k] = ex; ex += t[k]; }
	}
	__syncthreads();
#pragma unroll
	for (int i = 0; i < kScanIters; ++i) {
		const uint32_t b = (uint32_t)(i * kBucketThreads + tid);
		if (b < n) { const uint32_t ex = warp_tot[i * W + wid] + inc[i] - v[i]; a[b] = ex; cur[b] = ex; }
		if (crowd && v[i] > kSmallBin) { const uint32_t k = atomicAdd(&crowd[0], 1u); if (k < kMaxCrowded) crowd[1 + k] = b; }
	}
	__syncthreads();
}

// finishes bins [b0, b1) of a binned key array (wk[0] is the key at position base, count[b] = start of bin b): small bins by one
// thread each (insertion sort), crowded ones (listed in crowd[] on the way) by a whole warp afterwards.  Called by the whole CTA.
__device__ __forceinline__ void warp_sort_bin(uint64_t *w, uint32_t m, int lane)
{
	uint32_t pad = 32; while (pad < m) pad <<= 1;
	// bitonic network in its all-ascending form (first step of every merge compares mirrored positions), over a virtual
	// power-of-two range whose slots >= m hold +inf: every comparator moves the minimum to the lower index, so the
	// virtual keys never move and only pairs of real keys are touched
	for (uint32_t size = 2; size <= pad; size <<= 1) {
		const uint32_t half = size >> 1;
		for (uint32_t t = lane; t < (pad >> 1); t += 32) {
			const uint32_t blk = t / half, off = t - blk * half;
			const uint32_t l0 = blk * size + off, h0 = blk * size + size - 1u - off;
			if (h0 < m) { const uint64_t x = w[l0], y = w[h0]; if (x > y) { w[l0] = y; w[h0] = x; } }
		}
		__syncwarp();
		for (uint32_t stride = half >> 1; stride > 0; stride >>= 1) {
			for (uint32_t t = lane; t < (pad >> 1); t += 32) {
				const uint32_t l0 = 2 * t - (t & (stride - 1)), h0 = l0 + stride;
				if (h0 < m) { const uint64_t x = w[l0], y = w[h0]; if (x > y) { w[l0] = y; w[h0] = x; } }
			}
			__syncwarp();
		}
	}
}

__device__ __forceinline__ void finish_bins(uint64_t *wk, uint32_t base, const uint32_t *count, uint32_t b0, uint32_t b1, int tid, int lane, int wid,
                                            uint32_t *crowd)
{
	if (tid == 0) crowd[0] = 0;
	__syncthreads();
	for (uint32_t b = b0 + tid; b < b1; b += kBucketThreads) {
		const uint32_t lo = count[b], m = count[b + 1] - lo;
		if (m < 2) continue;
		if (m > kSmallBin) { const uint32_t k = atomicAdd(&crowd[0], 1u); if (k < kMaxCrowded) crowd[1 + k] = b; continue; }
		uint64_t *w = wk + (lo - base);
		for (uint32_t a = 1; a < m; ++a) {
			const uint64_t x = w[a];
			uint32_t c = a;
			while (c > 0 && w[c - 1] > x) { w[c] = w[c - 1]; --c; }
			w[c] = x;
		}
	}
	__syncthreads();
	const uint32_t nc = crowd[0];
	if (nc <= kMaxCrowded) {
		for (uint32_t k = wid; k < nc; k += kBucketThreads / 32) {  // warp-uniform loop
			const uint32_t b = crowd[1 + k], lo = count[b];
			warp_sort_bin(wk + (lo - base), count[b + 1] - lo, lane);
		}
	} else {                                                       // more crowded bins than the list holds: look at every bin
		for (uint32_t b = b0 + wid; b < b1; b += kBucketThreads / 32) {
			const uint32_t lo = count[b], m = count[b + 1] - lo;
			if (m > kSmallBin) warp_sort_bin(wk + (lo - base), m, lane);
		}
	}
}

// The bins of a binned key array straight to their sorted place in global memory: one thread per KEY -- its rank among the keys of
// its bin (ties by position, so the result is a permutation) is its offset in the bin.  src[0] is the key at position `base`,
// count[b] the start of bin b, bin_of(key) the key's bin; positions [p0, p1); dst is indexed by position.  Adjacent threads hold
// keys of the same or neighbouring bins, so the rank loops of a warp are about equally long (the one-thread-per-bin insertion
// sort this replaces ran 4 of 32 lanes wide).  Bins of more than kSmallBin keys are left to crowded_write.
template <class BinOf>
__device__ __forceinline__ void rank_write(const uint64_t *src, uint32_t base, const uint32_t *count, uint32_t p0, uint32_t p1, BinOf bin_of,
                                           uint64_t *__restrict__ dst, int tid)
{
	for (uint32_t p = p0 + tid; p < p1; p += kBucketThreads) {
		const uint64_t x = src[p - base];
		const uint32_t b = bin_of(x);
		const uint32_t lo = count[b], m = count[b + 1] - lo;
		if (m > kSmallBin) continue;
		const uint64_t *w = src + (lo - base);
		uint32_t r = 0;
		for (uint32_t q = 0; q < m; ++q) { const uint64_t y = w[q]; r += (y < x || (y == x && lo + q < p)) ? 1u : 0u; }
		dst[lo + r] = x;
	}
}

// the crowded bins: sorted in place by a whole warp each, then copied out.  crowd == nullptr or an overflowed list: every bin of
// [b0, b1) is looked at.  Warp-uniform.
__device__ __forceinline__ void crowded_write(uint64_t *src, uint32_t base, const uint32_t *count, uint32_t b0, uint32_t b1, const uint32_t *crowd,
                                              uint64_t *__restrict__ dst, int lane, int wid)
{
	auto one = [&](uint32_t b) {
		const uint32_t lo = count[b], m = count[b + 1] - lo;
		uint64_t *w = src + (lo - base);
		warp_sort_bin(w, m, lane);
		for (uint32_t t = lane; t < m; t += 32) dst[lo + t] = w[t];
	};
	const uint32_t nc = crowd ? crowd[0] : 0xffffffffu;
	if (nc <= kMaxCrowded) { for (uint32_t k = wid; k < nc; k += kBucketThreads / 32) one(crowd[1 + k]); }
	else for (uint32_t b = b0 + wid; b < b1; b += kBucketThreads / 32) if (count[b + 1] - count[b] > kSmallBin) one(b);
}

__device__ __forceinline__ uint32_t key_bin(uint64_t k, const uint32_t *__restrict__ bin_base, uint32_t tshift)
{
	return bin_base[key_list(k)] + (key_target(k) >> tshift);
}

template <uint32_t MAXK, uint32_t MINPAD>
__global__ void __launch_bounds__(kBucketThreads)
k_sort_buckets(uint64_t *__restrict__ keys, const uint64_t *__restrict__ seg_base, const uint32_t *__restrict__ seg_n, const uint32_t *__restrict__ seg_pad,
               const uint32_t *__restrict__ skip, uint32_t n_act, const uint32_t *__restrict__ bin_base, uint32_t tshift, uint32_t n_bins,
               uint64_t *__restrict__ stage /* MAXK == 0 only: arena parallel to keys */)
{
	extern __shared__ __align__(16) unsigned char bsm_raw[];
	BucketSmemT<MAXK> &sm = *reinterpret_cast<BucketSmemT<MAXK> *>(bsm_raw);
	const uint32_t i = blockIdx.x;
	if (i >= n_act || skip[i]) return;
	const uint32_t n = seg_n[i];
	if (n < 2 || (MAXK && n > MAXK) || n <= MINPAD) return;       // this instantiation's size class only (MINPAD < n <= MAXK; MAXK == 0: no upper limit)
	uint64_t *g = keys + seg_base[i];
	uint64_t *const wk = MAXK ? sm.keys : stage + seg_base[i];    // where the binned keys live while the bins are finished
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	for (uint32_t b = tid; b <= n_bins; b += kBucketThreads) sm.count[b] = 0;
	if (tid == 0) sm.crowd[0] = 0;
	__syncthreads();
	for (uint32_t k = tid; k < n; k += kBucketThreads) atomicAdd(&sm.count[key_bin(g[k], bin_base, tshift)], 1u);
	__syncthreads();
	// exclusive scan of count[0 .. n_bins) -> cursor / count (count[b] becomes the start of bin b, count[n_bins] = n)
	cta_exclusive_scan(sm.count, sm.cursor, n_bins, sm.warp_tot, tid, lane, wid, MAXK ? sm.crowd : nullptr);
	if (tid == 0) sm.count[n_bins] = n;
	__syncthreads();
	for (uint32_t k = tid; k < n; k += kBucketThreads) {
		const uint64_t key = g[k];
		wk[atomicAdd(&sm.cursor[key_bin(key, bin_base, tshift)], 1u)] = key;
	}
	__syncthreads();
	const auto bin1 = [&](uint64_t key) { return key_bin(key, bin_base, tshift); };
	if (MAXK) {                                                   // binned keys in shared memory -> sorted keys in global memory
		rank_write(wk, 0, sm.count, 0, n, bin1, g, tid);
		__syncthreads();
		crowded_write(wk, 0, sm.count, 0, n_bins, sm.crowd, g, lane, wid);
		return;
	}
	// MAXK == 0: the binned keys sit in global memory.  Runs of consecutive bins that fit the group buffer are loaded into shared
	// memory, binned a second time 2^F times finer (a long segment leaves tens of keys per first-level bin), finished there and
	// written straight to their final place; a single bin larger than the buffer is finished where it is.
	uint32_t b0 = 0;
	while (b0 < n_bins) {                                          // CTA-uniform
		const uint32_t lo = sm.count[b0];
		uint32_t b1 = b0, step = 1;
		auto fits = [&](uint32_t b) { return b <= n_bins && b - b0 <= kFineBins && sm.count[b] - lo <= kGroupKeys; };
		while (fits(b1 + step)) { b1 += step; step <<= 1; }        // gallop, then bisect
		for (step >>= 1; step > 0; step >>= 1) if (fits(b1 + step)) b1 += step;
		if (b1 == b0) {                                             // one crowded bin beyond the buffer
			b1 = b0 + 1;
			finish_bins(wk, 0, sm.count, b0, b1, tid, lane, wid, sm.crowd);
			__syncthreads();
			for (uint32_t k = lo + tid; k < sm.count[b1]; k += kBucketThreads) g[k] = wk[k];
		} else {
			const uint32_t m = sm.count[b1] - lo, nb = b1 - b0;
			uint32_t F = 0;
			while (F < tshift && (nb << (F + 1)) <= kFineBins && (m >> F) > 2u * nb) ++F;   // until ~2 keys per fine bin
			for (uint32_t k = tid; k < m; k += kBucketThreads) sm.keys[k] = wk[lo + k];
			if (F == 0) {
				__syncthreads();
				rank_write(sm.keys, lo, sm.count, lo, lo + m, bin1, g, tid);
				__syncthreads();
				crowded_write(sm.keys, lo, sm.count, b0, b1, nullptr, g, lane, wid);
			} else {
				const uint32_t nf = nb << F, fmask = (1u << F) - 1u, fsh = tshift - F;
				const auto bin2 = [&](uint64_t key) { return ((key_bin(key, bin_base, tshift) - b0) << F) | ((key_target(key) >> fsh) & fmask); };
				for (uint32_t b = tid; b <= nf; b += kBucketThreads) sm.fine[b] = 0;
				if (tid == 0) sm.crowd[0] = 0;
				__syncthreads();
				for (uint32_t k = tid; k < m; k += kBucketThreads) {
					const uint64_t key = sm.keys[k];
					atomicAdd(&sm.fine[bin2(key)], 1u);
				}
				__syncthreads();
				cta_exclusive_scan(sm.fine, sm.cursor, nf, sm.warp_tot, tid, lane, wid, sm.crowd);   // (cursor[] is free once the first-level scatter is done)
				if (tid == 0) sm.fine[nf] = m;
				__syncthreads();
				for (uint32_t k = tid; k < m; k += kBucketThreads) {
					const uint64_t key = sm.keys[k];
					sm.keys2[atomicAdd(&sm.cursor[bin2(key)], 1u)] = key;
				}
				__syncthreads();
				rank_write(sm.keys2, 0, sm.fine, 0, m, bin2, g + lo, tid);
				__syncthreads();
				crowded_write(sm.keys2, 0, sm.fine, 0, nf, sm.crowd, g + lo, lane, wid);
			}
		}
		__syncthreads();
		b0 = b1;
	}
}

} // namespace mapk
