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
//                below the threshold known so far); slots for their parts and tiles are reserved with
//                one atomicAdd per warp, the tile table is filled (no scan, no host round trip)
//   k_sp_parts   one CTA per tile of up to 256 consecutive parts of one chain: descriptors from the
//                anchors, the tile's read-event span and reference span staged in shared memory with
//                coalesced loads, parts counting-sorted by evaluator class and height so the lanes of
//                a warp run the same code equally long, one part per thread (register-resident
//                anti-diagonal evaluator for R' <= 4, shared-memory rings up to R' = 6, a whole warp
//                for anything wider); cost (and, when asked, cell count) per part to HBM
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
// One thread per group decides which of its chains are evaluated in this phase, reserves their slots in
// the part arrays and their tiles (one atomicAdd per warp), and fills the tile table.
__device__ __forceinline__ bool chain_active(const ChainLaunch &L, const SparsePlan &P, int phase, uint32_t g, uint32_t k, uint32_t k0,
                                             const ra_chain_t &ch)
{
	if (ch.n_anchors < 2) return false;
	if (phase == 1) return k == k0;                                           // threshold 0.0f: decided by the replay
	const float budget = chain_budget(L, ch);
	if (phase == 0) return !(budget < (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED));
	return (k != k0) && !(budget < P.best[g]);
}

__global__ void __launch_bounds__(256)
k_sp_prep(ChainLaunch L, SparsePlan P, int phase)
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	const int lane = threadIdx.x & 31;
	uint32_t k0 = 0, k1 = 0;
	if (g < L.n_groups) { k0 = L.group_first[g]; k1 = L.group_first[g + 1]; }
	uint32_t my_parts = 0, my_tiles = 0;
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		if (chain_active(L, P, phase, g, k, k0, ch)) { my_parts += ch.n_anchors - 1u; my_tiles += (ch.n_anchors - 1u + kTileParts - 1) / kTileParts; }
	}
	uint32_t ip = my_parts, it = my_tiles;                                    // inclusive warp scans
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t yp = __shfl_up_sync(0xffffffffu, ip, o), yt = __shfl_up_sync(0xffffffffu, it, o);
		if (lane >= o) { ip += yp; it += yt; }
	}
	unsigned long long bp = 0, bt = 0;
	if (lane == 31) { bp = atomicAdd(P.totals, (unsigned long long)ip); bt = atomicAdd(P.totals + 1, (unsigned long long)it); }
	bp = __shfl_sync(0xffffffffu, bp, 31); bt = __shfl_sync(0xffffffffu, bt, 31);
	uint64_t pb = bp + ip - my_parts, tb = bt + it - my_tiles;
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		const bool active = chain_active(L, P, phase, g, k, k0, ch);
		const uint32_t parts = active ? ch.n_anchors - 1u : 0u;
		P.n_parts[k] = parts;
		P.part_base[k] = pb;
		for (uint32_t t0 = 0; t0 < parts; t0 += kTileParts) { P.tile_chain[tb] = k; P.tile_t0[tb] = t0; ++tb; }
		pb += parts;
	}
}

struct SpShared {
	float sA[kStageCap];
	float sB[kStageCap];
	float rings[kSpWarps][2 * kNarrowRing * 32];
	uint64_t desc[kTileParts];                   // a_rel:11 | b_rel:11 | a_len:12 | b_len:12 | R' (0 = ring path):4 | band+1:8
	uint16_t order[kTileParts];                  // sorted narrow parts from the front, wide parts from the back
	uint32_t hist[kKeyBins];
	uint32_t n_big;
	uint32_t tile;
};

constexpr int kDiagMaxRp = 4;                    // register evaluator instantiated for R' = 1..4

template <bool COUNT>
__global__ void __launch_bounds__(kSpThreads)
k_sp_parts(ChainLaunch L, SparsePlan P)
{
	__shared__ SpShared sh;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	float *my_row = &sh.rings[wid][lane];
	float *my_bv = my_row + kNarrowRing * 32;
	float *my_gring = L.ring_scratch ? L.ring_scratch + (size_t)(blockIdx.x * kSpWarps + wid) * 3 * L.ring_cap : nullptr;
	const uint64_t n_tiles = P.totals[1];
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
		const uint32_t t0 = P.tile_t0[tile];
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
				const int rp = band >= 0 ? slant_radius(n, m, band) : 0;
				const bool diag = rp >= 1 && rp <= kDiagMaxRp;
				const bool narrow = staged && al < 4096u && bl < 4096u && (diag || narrow_ok(n, m, band));
				sh.desc[idx] = (uint64_t)(ps.query_position - q0) | ((uint64_t)(ps.target_position - r0) << 11) | ((uint64_t)(al & 4095u) << 22) |
				               ((uint64_t)(bl & 4095u) << 34) | ((uint64_t)(diag ? rp : 0) << 46) | ((uint64_t)(uint32_t)((band + 1) & 255) << 50);
				// sort key: evaluator class (R' = 1 | anything else) then height, so a warp runs one code path and equally long
				if (narrow) { key[u] = min(n, kKeyBins / 2 - 1) + (rp == 1 ? 0 : kKeyBins / 2); atomicAdd(&sh.hist[key[u]], 1u); }
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
			const uint64_t dsc = sh.desc[idx];
			const uint32_t al = (uint32_t)(dsc >> 22) & 4095u, bl = (uint32_t)(dsc >> 34) & 4095u;
			const float *a = sh.sA + ((uint32_t)dsc & 2047u), *b = sh.sB + ((uint32_t)(dsc >> 11) & 2047u);
			const bool swap = al < bl;                                           // dtw.cpp:284-292: A is the longer sequence
			const SmemLoader LA{swap ? b : a}, LB{swap ? a : b};
			const int n = (int)(swap ? bl : al), m = (int)(swap ? al : bl);
			uint32_t cnt = 0;
			float c;
			switch ((uint32_t)(dsc >> 46) & 15u) {
			case 1: c = thread_dtw_diag<1, COUNT>(LA, n, LB, m, &cnt); break;
			case 2: c = thread_dtw_diag<2, COUNT>(LA, n, LB, m, &cnt); break;
			case 3: c = thread_dtw_diag<3, COUNT>(LA, n, LB, m, &cnt); break;
			case 4: c = thread_dtw_diag<4, COUNT>(LA, n, LB, m, &cnt); break;
			default: c = thread_dtw_narrow(LA, n, LB, m, (int)((uint32_t)(dsc >> 50) & 255u) - 1, my_row, my_bv, &cnt); break;
			}
			if (t0 + idx != parts - 1) c -= fabsf(a[al - 1] - b[bl - 1]);         // exclude_last: rmap.cpp:270, dtw.cpp:514-516
			P.part_cost[out0 + idx] = c;
			if (COUNT) P.part_cells[out0 + idx] = cnt;
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
			if (lane == 0) { P.part_cost[out0 + idx] = c; if (COUNT) P.part_cells[out0 + idx] = (uint32_t)min(cnt, (uint64_t)0xffffffffu); cells_issued += cnt; }
			__syncwarp();
		}
	}
	if (COUNT && L.cells) {
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
			const float c = (base + lane < parts) ? pc[base + lane] : 0.0f;
			uint32_t done = cntb;
			const float cur0 = cur_max, cost0 = dtw_cost;
			bool hit = false;
			if (cntb == 32) {                                                    // full block: branch-free, same adds in the same order
#pragma unroll
				for (int t = 0; t < 32; ++t) {
					hit |= (cur_max < ms);                                       // :265-268 (the check before part base+t)
					const float sub = __shfl_sync(0xffffffffu, c, t);
					dtw_cost += sub;                                             // :279
					cur_max -= sub;                                              // :280
				}
			}
			if (cntb != 32 || hit) {                                             // tail block, or locate the part the reference stops at
				cur_max = cur0; dtw_cost = cost0;
				for (done = 0; done < cntb; ++done) {
					if (cur_max < ms) { dropped = true; break; }
					const float sub = __shfl_sync(0xffffffffu, c, (int)done);
					dtw_cost += sub;
					cur_max -= sub;
				}
			}
			if (L.cells) {                                                       // useful cells = cells of the parts the reference evaluates
				uint32_t ce = ((uint32_t)lane < done) ? pn[base + lane] : 0u;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) ce += __shfl_xor_sync(0xffffffffu, ce, o);
				useful += ce;
			}
		}
		// sum of read_region_size over all parts (:260, :292): anchors are shared by neighbouring parts
		n_aligned = parts ? (anc[0].query_position - anc[parts].query_position) + parts : 0u;
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
	cudaError_t e = cudaMemsetAsync(P.totals, 0, 2 * sizeof(unsigned long long), s);
	if (e != cudaSuccess) return e;
	k_sp_prep<<<(L.n_groups + 255) / 256, 256, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

cudaError_t launch_sparse_eval(const ChainLaunch &L, const SparsePlan &P, int phase, int n_sm, cudaStream_t s)
{
	cudaError_t e = cudaMemsetAsync(L.work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	if (L.cells) k_sp_parts<true><<<(uint32_t)n_sm * kSparseCtasPerSm, kSpThreads, 0, s>>>(L, P);
	else k_sp_parts<false><<<(uint32_t)n_sm * kSparseCtasPerSm, kSpThreads, 0, s>>>(L, P);
	k_sp_replay<<<(uint32_t)(((uint64_t)L.n_groups * 32 + 127) / 128), 128, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

} // namespace ra
