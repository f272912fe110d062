// chains_sparse.cu -- batched align_chain(cigar=false) for the SPARSE border constraint (the reference's
// default, src/rmap.cpp:238-300) plus the evaluation loop of gen_chains (:515-524).
//
// A sparse chain is a run of small, independent anchor-to-anchor DTW problems ("parts") laid end to
// end along the read and the reference.  Only the *replay* of a chain is sequential (cost += sub,
// budget -= sub, early exit, rmap.cpp:265-280) and only the threshold couples the chains of a group
// (min_score = best score so far, :515-524); the part costs themselves are independent.  So the work
// is split into a flat, fully parallel part-cost kernel and a tiny replay kernel:
//
//   k_sp_prep    per group: which chains can still matter (their budget chain_read_span*bonus is not
//                below the threshold known so far), parts / tiles per chain
//   (scan x2)    part_base[], tile_base[]
//   k_sp_tiles   tile -> chain table
//   k_sp_parts   one CTA per tile of up to 256 consecutive parts of one chain: descriptors from the
//                anchors, the tile's read-event span and reference span staged in shared memory with
//                coalesced loads, parts counting-sorted by height so the lanes of a warp run equally
//                long, one part per thread (register-window row sweep for R' <= 2, shared-memory rings
//                up to R' = 6, a whole warp for anything wider); cost and cell count per part to HBM
//   k_sp_replay  one warp per group: the reference's serial float arithmetic in part order (every lane
//                carries the same state, values broadcast by shuffle), alignment_score (:306), keep
//                flag (:518), running best (:519-521)
//
// With the running threshold (min_score_mode 1) the pipeline runs twice: phase A evaluates the first
// chain of every group, phase B the remaining chains whose budget is not already below the best score
// of phase A (a superset of what the reference evaluates; the replay applies the exact rule, so
// scores and keep flags are the reference's -- only cells_issued can exceed cells_useful).
#include "dtw_core.cuh"

namespace ra {

constexpr int kSpThreads = 128;
constexpr int kSpWarps = kSpThreads / 32;
constexpr int kTileParts = 256;                  // two parts per thread
constexpr int kStageCap = 2048;                  // floats of read span / reference span staged per tile
constexpr uint32_t kSpWarpRingCap = 256;         // 3 rings of 256 floats alias the warp's 2*16*32-float lane rings
constexpr int kKeyBins = 64;

__device__ __forceinline__ int part_band(const ChainLaunch &L, uint32_t a_len)
{
	return (L.fill == RA_DTW_FILL_METHOD_BANDED) ? max(1, (int)((float)a_len * L.band_frac)) : -1;   // rmap.cpp:276
}

__device__ __forceinline__ float chain_budget(const ChainLaunch &L, const ra_chain_t &ch)
{
	const ra_anchor_t *anc = L.anchors + ch.anchor_off;
	const uint32_t span = anc[0].query_position - anc[ch.n_anchors - 1].query_position + 1;            // :245
	return (float)span * L.match_bonus;                                                                 // :246
}

// phase 0: min_score_mode 0 (every chain against its own min_score); phase 1 / 2: A / B of mode 1.
__global__ void __launch_bounds__(256)
k_sp_prep(ChainLaunch L, SparsePlan P, int phase)
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= L.n_groups) return;
	const uint32_t k0 = L.group_first[g], k1 = L.group_first[g + 1];
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		const uint32_t parts = ch.n_anchors - 1u;
		bool active = parts > 0;
		if (active) {
			const float budget = chain_budget(L, ch);
			if (phase == 0) active = !(budget < (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED));
			else if (phase == 1) active = (k == k0);                       // threshold 0.0f: decided by the replay
			else active = (k != k0) && !(budget < P.best[g]);
		}
		P.n_parts[k] = active ? parts : 0u;
		P.n_tiles[k] = active ? (parts + kTileParts - 1) / kTileParts : 0u;
	}
}

__global__ void __launch_bounds__(256)
k_sp_tiles(uint32_t n_chains, SparsePlan P)
{
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n_chains) return;
	const uint64_t b = P.tile_base[k], e = P.tile_base[k + 1];
	for (uint64_t t = b; t < e; ++t) P.tile_chain[t] = k;
}

struct SpShared {
	float sA[kStageCap];
	float sB[kStageCap];
	float rings[kSpWarps][2 * kNarrowRing * 32];
	uint16_t a_rel[kTileParts], b_rel[kTileParts], a_len[kTileParts], b_len[kTileParts];
	uint16_t order[kTileParts];                  // sorted narrow parts from the front, wide parts from the back
	uint32_t hist[kKeyBins];
	uint32_t n_big;
	uint32_t tile;
};

__global__ void __launch_bounds__(kSpThreads)
k_sp_parts(ChainLaunch L, SparsePlan P)
{
	__shared__ SpShared sh;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	float *my_row = &sh.rings[wid][lane];
	float *my_bv = my_row + kNarrowRing * 32;
	float *my_gring = L.ring_scratch ? L.ring_scratch + (size_t)(blockIdx.x * kSpWarps + wid) * 3 * L.ring_cap : nullptr;
	const uint64_t n_tiles = P.tile_base[P.n_chains];
	uint64_t cells_issued = 0;

	for (;;) {
		__syncthreads();                                                     // previous tile fully consumed
		if (tid == 0) sh.tile = atomicAdd(L.work_counter, 1u);
		if (tid < kKeyBins) sh.hist[tid] = 0;
		if (tid == 0) sh.n_big = 0;
		__syncthreads();
		const uint32_t tile = sh.tile;
		if (tile >= n_tiles) break;
		const uint32_t k = P.tile_chain[tile];
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *__restrict__ anc = L.anchors + ch.anchor_off;
		const float *__restrict__ read = L.read_arena + ch.read_off;
		const float *__restrict__ ref = L.ref_arena + ch.ref_off;
		const uint32_t parts = ch.n_anchors - 1u;
		const uint32_t t0 = (uint32_t)(tile - P.tile_base[k]) * kTileParts;
		const uint32_t tn = min((uint32_t)kTileParts, parts - t0);
		const uint64_t out0 = P.part_base[k] + t0;
		// ---- 1. descriptors (sparse part p runs from anchors[parts-p] to anchors[parts-p-1], rmap.cpp:253-260) ----
		const ra_anchor_t s0 = anc[parts - t0];
		const uint32_t q0 = s0.query_position, r0 = s0.target_position;
		int key[2] = { -1, -1 };
#pragma unroll
		for (int u = 0; u < 2; ++u) {
			const uint32_t idx = (uint32_t)tid + (uint32_t)u * kSpThreads;
			if (idx < tn) {
				const ra_anchor_t ps = anc[parts - (t0 + idx)], pe = anc[parts - (t0 + idx) - 1];
				const uint32_t al = pe.query_position - ps.query_position + 1;         // :260
				const uint32_t bl = pe.target_position - ps.target_position + 1;       // :257
				const int n = (int)max(al, bl), m = (int)min(al, bl);
				const int band = part_band(L, al);
				// staged = both spans of the part lie inside the tile's staging window
				const bool staged = (pe.query_position - q0 < (uint32_t)kStageCap) && (pe.target_position - r0 < (uint32_t)kStageCap);
				const bool narrow = staged && (win6_ok(n, m, band) || narrow_ok(n, m, band));
				sh.a_rel[idx] = (uint16_t)(ps.query_position - q0); sh.b_rel[idx] = (uint16_t)(ps.target_position - r0);
				sh.a_len[idx] = (uint16_t)min(al, 65535u); sh.b_len[idx] = (uint16_t)min(bl, 65535u);
				if (narrow) { key[u] = min(n, kKeyBins - 1); atomicAdd(&sh.hist[key[u]], 1u); }
				else { const uint32_t b = atomicAdd(&sh.n_big, 1u); sh.order[kTileParts - 1 - b] = (uint16_t)idx; }
			}
		}
		// ---- 2. stage the tile's spans (clipped to the staging window) ----
		{
			const ra_anchor_t e_last = anc[parts - (t0 + tn - 1) - 1];
			const uint32_t la = min(e_last.query_position - q0 + 1, (uint32_t)kStageCap);
			const uint32_t lb = min(e_last.target_position - r0 + 1, (uint32_t)kStageCap);
			for (uint32_t x = tid; x < la; x += kSpThreads) sh.sA[x] = __ldg(read + q0 + x);
			for (uint32_t x = tid; x < lb; x += kSpThreads) sh.sB[x] = __ldg(ref + r0 + x);
		}
		__syncthreads();
		// ---- 3. exclusive scan of the histogram (warp 0, two bins per lane), scatter ----
		if (wid == 0) {
			const uint32_t v0 = sh.hist[2 * lane], v1 = sh.hist[2 * lane + 1];
			uint32_t incl = v0 + v1;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
			const uint32_t ex = incl - (v0 + v1);
			sh.hist[2 * lane] = ex; sh.hist[2 * lane + 1] = ex + v0;
		}
		__syncthreads();
#pragma unroll
		for (int u = 0; u < 2; ++u)
			if (key[u] >= 0) { const uint32_t pos = atomicAdd(&sh.hist[key[u]], 1u); sh.order[pos] = (uint16_t)((uint32_t)tid + (uint32_t)u * kSpThreads); }
		__syncthreads();
		const uint32_t n_big = sh.n_big, n_narrow = tn - n_big;
		// ---- 4a. narrow parts: one per thread; ranks tid and (n_narrow-1-tid) pair a short with a tall part ----
#pragma unroll 1
		for (int u = 0; u < 2; ++u) {
			const int pos = (u == 0) ? tid : (int)n_narrow - 1 - tid;
			if (pos < 0 || pos >= (int)n_narrow || (u == 1 && pos < kSpThreads)) continue;
			const uint32_t idx = sh.order[pos];
			const uint32_t al = sh.a_len[idx], bl = sh.b_len[idx];
			const float *a = sh.sA + sh.a_rel[idx], *b = sh.sB + sh.b_rel[idx];
			const int band = part_band(L, al);
			const bool swap = al < bl;                                           // dtw.cpp:284-292: A is the longer sequence
			const SmemLoader LA{swap ? b : a}, LB{swap ? a : b};
			const int n = (int)(swap ? bl : al), m = (int)(swap ? al : bl);
			uint32_t cnt = 0;
			float c;
			if (win6_ok(n, m, band)) c = thread_dtw_win6(LA, n, LB, m, band, &cnt);
			else c = thread_dtw_narrow(LA, n, LB, m, band, my_row, my_bv, &cnt);
			if (t0 + idx != parts - 1) c -= fabsf(a[al - 1] - b[bl - 1]);         // exclude_last: rmap.cpp:270, dtw.cpp:514-516
			P.part_cost[out0 + idx] = c; P.part_cells[out0 + idx] = cnt;
			cells_issued += cnt;
		}
		// ---- 4b. wide parts: one warp each, operands from global memory ----
		__syncwarp();
		for (uint32_t bk = wid; bk < n_big; bk += kSpWarps) {
			const uint32_t idx = sh.order[kTileParts - 1 - bk];
			const uint32_t p = t0 + idx;
			const ra_anchor_t s = anc[parts - p], e = anc[parts - p - 1];
			const uint32_t al = e.query_position - s.query_position + 1, bl = e.target_position - s.target_position + 1;
			const float *a = read + s.query_position, *b = ref + s.target_position;
			const int band = part_band(L, al);
			const float *A = a, *B = b; int n = (int)al, m = (int)bl;
			if (al < bl) { A = b; B = a; n = (int)bl; m = (int)al; }
			const int width = band_width_of(n, m, band);
			uint32_t cap = 1; while ((int)cap < width) cap <<= 1;
			float *ring = &sh.rings[wid][0];
			bool ok = true;
			if (cap > kSpWarpRingCap) {
				if (!my_gring || cap > L.ring_cap) { ok = false; if (lane == 0) atomicMax(L.need_cap, cap); }
				else { ring = my_gring; cap = L.ring_cap; }
			} else cap = kSpWarpRingCap;
			uint64_t cnt = 0; float c = 0.0f;
			if (ok) {
				c = warp_dtw_cost(A, n, B, m, band, ring, cap, lane, &cnt);
				if (p != parts - 1) c -= fabsf(__ldg(A + n - 1) - __ldg(B + m - 1));
			}
			if (lane == 0) { P.part_cost[out0 + idx] = c; P.part_cells[out0 + idx] = (uint32_t)min(cnt, (uint64_t)0xffffffffu); cells_issued += cnt; }
			__syncwarp();
		}
	}
	if (L.cells) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) cells_issued += __shfl_down_sync(0xffffffffu, cells_issued, o);
		if (lane == 0 && cells_issued) atomicAdd(reinterpret_cast<unsigned long long *>(L.cells), (unsigned long long)cells_issued);
	}
}

// One warp per group; every lane carries the same replay state.
__global__ void __launch_bounds__(128)
k_sp_replay(ChainLaunch L, SparsePlan P, int phase)
{
	const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (g >= L.n_groups) return;
	const uint32_t k0 = L.group_first[g], k1 = L.group_first[g + 1];
	const uint32_t kb = (phase == 2) ? k0 + 1 : k0, ke = (phase == 1) ? min(k0 + 1, k1) : k1;
	float best = (phase == 2) ? P.best[g] : 0.0f;                               // rmap.cpp:515
	uint64_t useful = 0;
	for (uint32_t k = kb; k < ke; ++k) {
		const ra_chain_t ch = L.chains[k];
		const uint32_t parts = ch.n_anchors - 1u;
		const float ms = (L.min_score_mode == 1) ? best : (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED);
		float cur_max = chain_budget(L, ch);                                     // :246
		float dtw_cost = 0.0f;
		uint32_t n_aligned = 0;
		bool dropped = false;
		if (parts > 0 && P.n_parts[k] == 0) dropped = true;                      // budget already below the threshold: the check before part 0 fires
		const float *__restrict__ pc = P.part_cost + P.part_base[k];
		const uint32_t *__restrict__ pn = P.part_cells + P.part_base[k];
		const ra_anchor_t *__restrict__ anc = L.anchors + ch.anchor_off;
		for (uint32_t base = 0; base < parts && !dropped; base += 32) {
			const uint32_t cntb = min(32u, parts - base);
			const bool mine = base + lane < parts;
			const float c = mine ? pc[base + lane] : 0.0f;
			uint32_t done = 0;
			for (; done < cntb; ++done) {
				if (cur_max < ms) { dropped = true; break; }                     // :265-268
				const float sub = __shfl_sync(0xffffffffu, c, (int)done);
				dtw_cost += sub;                                                 // :279
				cur_max -= sub;                                                  // :280
			}
			// integer side for the parts consumed: read_region_size (:260, :292) and useful cells
			uint32_t al = 0, ce = 0;
			if ((uint32_t)lane < done) {
				const uint32_t p = base + lane;
				al = anc[parts - p - 1].query_position - anc[parts - p].query_position + 1;
				ce = pn[p];
			}
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) { al += __shfl_xor_sync(0xffffffffu, al, o); ce += __shfl_xor_sync(0xffffffffu, ce, o); }
			n_aligned += al; useful += ce;
		}
		float score = RA_SCORE_DROPPED;
		if (!dropped)                                                            // :306; the reference build fuses it when fp_contract is set
			score = L.fp_contract ? __fmaf_rn((float)n_aligned, L.match_bonus, -dtw_cost)
			                      : __fsub_rn(__fmul_rn((float)n_aligned, L.match_bonus), dtw_cost);
		const bool kp = score >= L.min_dtw_score;                                // :518
		if (lane == 0) { L.score[k] = score; if (L.keep) L.keep[k] = kp ? 1 : 0; }
		if (kp && score > best) best = score;                                    // :519-521
	}
	if (lane == 0) {
		if (phase == 1) P.best[g] = best;
		if (L.cells && useful) atomicAdd(reinterpret_cast<unsigned long long *>(L.cells + 1), (unsigned long long)useful);
	}
}

cudaError_t launch_sparse_prep(const ChainLaunch &L, const SparsePlan &P, int phase, cudaStream_t s)
{
	k_sp_prep<<<(L.n_groups + 255) / 256, 256, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

cudaError_t launch_sparse_eval(const ChainLaunch &L, const SparsePlan &P, int phase, int n_sm, cudaStream_t s)
{
	cudaError_t e = cudaMemsetAsync(L.work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	k_sp_tiles<<<(P.n_chains + 255) / 256, 256, 0, s>>>(P.n_chains, P);
	k_sp_parts<<<(uint32_t)n_sm * kSparseCtasPerSm, kSpThreads, 0, s>>>(L, P);
	k_sp_replay<<<(uint32_t)(((uint64_t)L.n_groups * 32 + 127) / 128), 128, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

} // namespace ra
