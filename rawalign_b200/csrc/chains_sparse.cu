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
	const bool invalid = *reinterpret_cast<volatile uint32_t *>(L.error) != 0;   // k_validate_chains found a malformed chain: plan no parts at all
	uint32_t my_parts = 0, my_tiles = 0;
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		if (!invalid && chain_active(L, P, phase, g, k, k0, ch)) { my_parts += ch.n_anchors - 1u; my_tiles += (ch.n_anchors - 1u + P.tile_parts - 1) / P.tile_parts; }
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
		const bool active = !invalid && chain_active(L, P, phase, g, k, k0, ch);
		const uint32_t parts = active ? ch.n_anchors - 1u : 0u;
		P.n_parts[k] = parts;
		P.part_base[k] = pb;
		for (uint32_t t0 = 0; t0 < parts; t0 += P.tile_parts) { P.tile_chain[tb] = k; P.tile_t0[tb] = t0; ++tb; }
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

// ---------------------------------------------------------------------------------------------------
// k_sp_parts2 (default): the same work as k_sp_parts with WARP-autonomous tiles.  Chains hold ~170 parts of ~3-13 events
// each; a CTA-wide tile of 256 left a third of the thread slots empty and made every warp wait at the CTA barriers for
// the slowest one (ncu on k_sp_parts: 15.6 of 32 lanes active, barrier the top stall reason, and the handful of R' >= 3
// parts of a tile -- 2 % of the parts -- cost 30 % of the warp instructions because they ran 1-3 lanes wide).  Here a
// warp owns a tile of up to 128 consecutive parts of one chain, synchronises only with itself, and moves on to its next
// tile as soon as it is done:
//   * the tile is cut into batches whose read / reference spans fit the warp's staging window (kSp2Stage floats each);
//   * R' <= 2 parts (98 % at banded=0.1: band 1 -> R' = 1 when square, 2 otherwise) are counting-sorted by height and
//     run 32 at a time through ONE register evaluator (R' = 1 embedded in the R' = 2 schedule, dtw_core.cuh), so a round
//     has no class divergence; a round of only R' = 1 parts takes the smaller R' = 1 evaluator;
//   * other narrow parts run in full rounds of 32 through the shared-memory ring evaluator; a short remainder is appended
//     to a global list and evaluated one part per thread by k_sp_rest, instead of occupying a warp 1-3 lanes wide;
//   * wide parts (band wider than the lane rings, or a part larger than the staging window): the whole warp, operands
//     from global memory, as before.
// ---------------------------------------------------------------------------------------------------
// |cell set| of one banded problem with slant-extended radius rp (dtw.cpp:352-366, :416-420; same walk as ra_dtw_cells)
__device__ __forceinline__ uint32_t diag_cells(int n, int m, int rp)
{
	const int P = rp + ((rp & 1) ? 0 : 1), S = rp + ((rp & 1) ? 1 : 0);
	auto span = [&](int start_i, int start_j, int len) {
		const int o0 = max(max(0, start_i - (n - 1)), -start_j), o1 = min(min(len, start_i + 1), m - start_j);
		return (uint32_t)max(0, o1 - o0);
	};
	uint32_t cells = 1;
	int r = 0, e = 0;
	for (int c = 1; c < n; ++c) {
		e += m;
		if (e >= n) { e -= n; ++r; cells += span(c + S / 2 - 1, r - S / 2, S); }
		cells += span(c + P / 2, r - P / 2, P);
	}
	return cells;
}

constexpr int kSp2Warps = 4;
constexpr int kSp2Tile = kSparseTileParts2;       // parts per warp tile
constexpr int kSp2Per = kSp2Tile / 32;
#ifndef RA_SP2_STAGE
#define RA_SP2_STAGE 480
#endif
constexpr int kSp2Stage = RA_SP2_STAGE;            // floats of read span / reference span staged per warp batch (<= 1024: 10-bit offsets in desc)
static_assert(kSp2Stage <= 1024 && kSp2Stage % 32 == 0, "staging window");
constexpr int kSp2Bins = 64;                      // class (0: R' <= 2, 1: other narrow) x height (32 bins)
constexpr uint32_t kSp2DeferBelow = 16;           // a remainder of fewer other-narrow parts goes to k_sp_rest

struct __align__(16) Sp2Warp {
	float sA[kSp2Stage];
	float sB[kSp2Stage];
	float rings[2 * kNarrowRing * 32];            // lane rings of thread_dtw_narrow; 3 x 256 floats of it double as the wide parts' warp ring
	uint64_t desc[kSp2Tile];                      // a_rel:10 | b_rel:10 | a_len:11 | b_len:11 | R':4 | band+1:8
	uint32_t hist[kSp2Bins];
	uint32_t n_wide;
	uint8_t order[kSp2Tile];                      // sorted narrow parts from the front, wide parts from the back
};
static_assert(2 * kNarrowRing * 32 >= 3 * (int)kSpWarpRingCap, "warp ring aliases the lane rings");

template <bool COUNT>
__global__ void __launch_bounds__(kSp2Warps * 32)
k_sp_parts2(ChainLaunch L, SparsePlan P)
{
	extern __shared__ __align__(16) unsigned char sp2_smem[];
	Sp2Warp &W = reinterpret_cast<Sp2Warp *>(sp2_smem)[threadIdx.x >> 5];
	constexpr uint32_t kFull = 0xffffffffu;
	const int lane = threadIdx.x & 31;
	const uint32_t gwarp = blockIdx.x * kSp2Warps + (threadIdx.x >> 5);
	float *my_row = &W.rings[lane];
	float *my_bv = my_row + kNarrowRing * 32;
	float *my_gring = L.ring_scratch ? L.ring_scratch + (size_t)gwarp * 3 * L.ring_cap : nullptr;
	const uint64_t n_tiles = P.totals[1];
	uint64_t cells_issued = 0;

	for (;;) {
		uint32_t tile = 0;
		if (lane == 0) tile = atomicAdd(L.work_counter, 1u);
		tile = __shfl_sync(kFull, tile, 0);
		if (tile >= n_tiles) break;
		const uint32_t k = P.tile_chain[tile];
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *__restrict__ anc = L.anchors + ch.anchor_off;
		const float *__restrict__ read = L.read_arena + ch.read_off;
		const float *__restrict__ ref = L.ref_arena + ch.ref_off;
		const uint32_t parts = ch.n_anchors - 1u;
		const uint32_t t0 = P.tile_t0[tile];
		const uint32_t tn = min((uint32_t)kSp2Tile, parts - t0);
		const uint64_t out0 = P.part_base[k] + t0;

		// one part by the whole warp, operands from global memory (tile-relative index idx)
		auto wide_part = [&](uint32_t idx) {
			const uint32_t p = t0 + idx;
			const ra_anchor_t s = anc[parts - p], e = anc[parts - p - 1];
			const uint32_t al = e.query_position - s.query_position + 1, bl = e.target_position - s.target_position + 1;
			const float *a = read + s.query_position, *b = ref + s.target_position;
			const int band = part_band(L, al);
			const float *A = a, *B = b; int n = (int)al, m = (int)bl;
			if (al < bl) { A = b; B = a; n = (int)bl; m = (int)al; }
			const int width = band_width_of(n, m, band);
			uint32_t cap = 1; while ((int)cap < width) cap <<= 1;
			float *ring = W.rings;
			bool ok = true;
			if (cap > kSpWarpRingCap) {
				if (!my_gring || cap > L.ring_cap) { ok = false; if (lane == 0) atomicMax(L.need_cap, cap); }
				else { ring = my_gring; cap = L.ring_cap; }
			} else cap = kSpWarpRingCap;
			uint64_t cnt = 0; float c = 0.0f;
			if (ok) {
				c = warp_dtw_cost(A, n, B, m, band, ring, cap, lane, &cnt);
				if (p != parts - 1) c -= fabsf(__ldg(A + n - 1) - __ldg(B + m - 1));     // exclude_last: rmap.cpp:270, dtw.cpp:514-516
			}
			if (lane == 0) { P.part_cost[out0 + idx] = c; if (COUNT) P.part_cells[out0 + idx] = (uint32_t)min(cnt, (uint64_t)0xffffffffu); cells_issued += cnt; }
			__syncwarp();
		};

		uint32_t done = 0;
		while (done < tn) {
			// ---- 1. descriptors of the batch: the parts from `done` on whose spans end inside the staging window ----
			const ra_anchor_t s0 = anc[parts - (t0 + done)];
			const uint32_t q0 = s0.query_position, r0 = s0.target_position;
			W.hist[lane] = 0; W.hist[lane + 32] = 0;
			if (lane == 0) W.n_wide = 0;
			__syncwarp();
			int key[kSp2Per];
			uint32_t bn = 0;                                                      // batch size (anchors are increasing, so "fits" is a prefix property)
			uint32_t la = 0, lb = 0;                                               // staged span lengths
#pragma unroll
			for (int u = 0; u < kSp2Per; ++u) {
				key[u] = -1;
				const uint32_t bi = (uint32_t)lane + 32u * (uint32_t)u;             // batch-relative index
				bool fits = false;
				if (done + bi < tn) {
					const uint32_t p = t0 + done + bi;
					const ra_anchor_t ps = anc[parts - p], pe = anc[parts - p - 1];
					fits = (pe.query_position - q0 < (uint32_t)kSp2Stage) && (pe.target_position - r0 < (uint32_t)kSp2Stage);
					if (fits) {
						const uint32_t al = pe.query_position - ps.query_position + 1;         // :260
						const uint32_t bl = pe.target_position - ps.target_position + 1;       // :257
						const int n = (int)max(al, bl), m = (int)min(al, bl);
						const int band = part_band(L, al);
						const int rp = band >= 0 ? slant_radius(n, m, band) : 0;
						const bool c0 = rp == 1 || rp == 2;
						const bool nar = c0 || narrow_ok(n, m, band);
						W.desc[bi] = (uint64_t)(ps.query_position - q0) | ((uint64_t)(ps.target_position - r0) << 10) | ((uint64_t)al << 20) |
						             ((uint64_t)bl << 31) | ((uint64_t)(uint32_t)min(rp, 15) << 42) | ((uint64_t)(uint32_t)((band + 1) & 255) << 46);
						if (nar) { key[u] = min(n, 31) + (c0 ? 0 : 32); atomicAdd(&W.hist[key[u]], 1u); }
						else { const uint32_t b = atomicAdd(&W.n_wide, 1u); W.order[kSp2Tile - 1 - b] = (uint8_t)bi; }
						la = max(la, pe.query_position - q0 + 1); lb = max(lb, pe.target_position - r0 + 1);
					}
				}
				bn += (uint32_t)__popc(__ballot_sync(kFull, fits));
			}
			if (bn == 0) { wide_part(done); done += 1; continue; }                 // a single part larger than the window
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) { la = max(la, __shfl_xor_sync(kFull, la, o)); lb = max(lb, __shfl_xor_sync(kFull, lb, o)); }
			// ---- 2. stage the batch's spans ----
			for (uint32_t x = lane; x < la; x += 32) W.sA[x] = __ldg(read + q0 + x);
			for (uint32_t x = lane; x < lb; x += 32) W.sB[x] = __ldg(ref + r0 + x);
			__syncwarp();
			// ---- 3. exclusive scan of the histogram (two bins per lane), scatter ----
			uint32_t n0;
			{
				const uint32_t v0 = W.hist[2 * lane], v1 = W.hist[2 * lane + 1];
				uint32_t incl = v0 + v1;
#pragma unroll
				for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(kFull, incl, o); if (lane >= o) incl += y; }
				const uint32_t ex = incl - (v0 + v1);
				n0 = __shfl_sync(kFull, ex, 16);                                    // bins 0..31 = class 0
				__syncwarp();
				W.hist[2 * lane] = ex; W.hist[2 * lane + 1] = ex + v0;
			}
			__syncwarp();
#pragma unroll
			for (int u = 0; u < kSp2Per; ++u)
				if (key[u] >= 0) { const uint32_t pos = atomicAdd(&W.hist[key[u]], 1u); W.order[pos] = (uint8_t)((uint32_t)lane + 32u * (uint32_t)u); }
			__syncwarp();
			const uint32_t n_wide = W.n_wide, n_narrow = bn - n_wide, n1 = n_narrow - n0;
			const uint32_t tb = done;                                              // tile-relative index of the batch's first part

			auto finish = [&](uint32_t bi, uint64_t dsc, float c, uint32_t cnt) {
				const uint32_t al = (uint32_t)(dsc >> 20) & 2047u, bl = (uint32_t)(dsc >> 31) & 2047u;
				const float *a = W.sA + ((uint32_t)dsc & 1023u), *b = W.sB + ((uint32_t)(dsc >> 10) & 1023u);
				if (t0 + tb + bi != parts - 1) c -= fabsf(a[al - 1] - b[bl - 1]);    // exclude_last: rmap.cpp:270, dtw.cpp:514-516
				P.part_cost[out0 + tb + bi] = c;
				if (COUNT) P.part_cells[out0 + tb + bi] = cnt;
				cells_issued += cnt;
			};
			// ---- 4a. R' <= 2: rounds of 32 equally tall parts through one register evaluator ----
			for (uint32_t base = 0; base < n0; base += 32) {
				const uint32_t pos = base + (uint32_t)lane;
				const bool active = pos < n0;
				uint32_t bi = 0; uint64_t dsc = 0;
				if (active) { bi = W.order[pos]; dsc = W.desc[bi]; }
				const uint32_t rp = (uint32_t)(dsc >> 42) & 15u;
				const bool any2 = __any_sync(kFull, active && rp == 2u);
				if (active) {
					const uint32_t al = (uint32_t)(dsc >> 20) & 2047u, bl = (uint32_t)(dsc >> 31) & 2047u;
					const float *a = W.sA + ((uint32_t)dsc & 1023u), *b = W.sB + ((uint32_t)(dsc >> 10) & 1023u);
					const bool swap = al < bl;                                       // dtw.cpp:284-292: A is the longer sequence
					const SmemLoader LA{swap ? b : a}, LB{swap ? a : b};
					const int n = (int)(swap ? bl : al), m = (int)(swap ? al : bl);
					float c;
					if (any2) c = thread_dtw_diag<2, false, true>(LA, n, LB, m, nullptr, rp == 1u);
					else c = thread_dtw_diag<1, false>(LA, n, LB, m, nullptr);
					finish(bi, dsc, c, COUNT ? diag_cells(n, m, (int)rp) : 0u);   // instrumentation stays out of the evaluator
				}
				__syncwarp();
			}
			// ---- 4b. other narrow parts: full rounds here, a short remainder to k_sp_rest ----
			{
				uint32_t base = 0;
				for (; base < n1; base += 32) {
					const uint32_t left = n1 - base;
					if (left < kSp2DeferBelow && P.rest) break;
					const uint32_t pos = n0 + base + (uint32_t)lane;
					if (pos < n_narrow) {
						const uint32_t bi = W.order[pos];
						const uint64_t dsc = W.desc[bi];
						const uint32_t al = (uint32_t)(dsc >> 20) & 2047u, bl = (uint32_t)(dsc >> 31) & 2047u;
						const float *a = W.sA + ((uint32_t)dsc & 1023u), *b = W.sB + ((uint32_t)(dsc >> 10) & 1023u);
						const bool swap = al < bl;
						const SmemLoader LA{swap ? b : a}, LB{swap ? a : b};
						const int n = (int)(swap ? bl : al), m = (int)(swap ? al : bl);
						uint32_t cnt = 0;
						const float c = thread_dtw_narrow(LA, n, LB, m, (int)((uint32_t)(dsc >> 46) & 255u) - 1, my_row, my_bv, &cnt);
						finish(bi, dsc, c, cnt);
					}
					__syncwarp();
				}
				if (base < n1) {                                                   // fewer than kSp2DeferBelow parts left
					const uint32_t cntd = n1 - base;
					unsigned long long slot = 0;
					if (lane == 0) slot = atomicAdd(P.totals + 2, (unsigned long long)cntd);
					slot = __shfl_sync(kFull, slot, 0);
					if ((uint32_t)lane < cntd) P.rest[slot + lane] = ((uint64_t)k << 32) | (uint64_t)(t0 + tb + W.order[n0 + base + lane]);
				}
			}
			// ---- 4c. wide parts ----
			__syncwarp();
			for (uint32_t bk = 0; bk < n_wide; ++bk) wide_part(tb + W.order[kSp2Tile - 1 - bk]);
			__syncwarp();
			done += bn;
		}
	}
	if (COUNT && L.cells) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) cells_issued += __shfl_down_sync(0xffffffffu, cells_issued, o);
		if (lane == 0 && cells_issued) atomicAdd(reinterpret_cast<unsigned long long *>(L.cells), (unsigned long long)cells_issued);
	}
}

// The deferred other-narrow parts, one per thread, operands through the read-only path.
template <bool COUNT>
__global__ void __launch_bounds__(128)
k_sp_rest(ChainLaunch L, SparsePlan P)
{
	__shared__ float rings[4][2 * kNarrowRing * 32];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	float *my_row = &rings[wid][lane];
	float *my_bv = my_row + kNarrowRing * 32;
	const uint64_t n_rest = P.totals[2];
	uint64_t cells_issued = 0;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rest; i += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t ent = P.rest[i];
		const uint32_t k = (uint32_t)(ent >> 32), p = (uint32_t)ent;
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *__restrict__ anc = L.anchors + ch.anchor_off;
		const uint32_t parts = ch.n_anchors - 1u;
		const ra_anchor_t s = anc[parts - p], e = anc[parts - p - 1];
		const uint32_t al = e.query_position - s.query_position + 1, bl = e.target_position - s.target_position + 1;
		const float *a = L.read_arena + ch.read_off + s.query_position, *b = L.ref_arena + ch.ref_off + s.target_position;
		const bool swap = al < bl;
		const LdgLoader LA{swap ? b : a}, LB{swap ? a : b};
		const int n = (int)(swap ? bl : al), m = (int)(swap ? al : bl);
		uint32_t cnt = 0;
		float c = thread_dtw_narrow(LA, n, LB, m, part_band(L, al), my_row, my_bv, &cnt);
		if (p != parts - 1) c -= fabsf(__ldg(a + al - 1) - __ldg(b + bl - 1));
		P.part_cost[P.part_base[k] + p] = c;
		if (COUNT) P.part_cells[P.part_base[k] + p] = cnt;
		cells_issued += cnt;
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
	const bool invalid = *reinterpret_cast<volatile uint32_t *>(L.error) != 0;
	for (uint32_t k = kb; k < ke; ++k) {
		if (invalid) { if (lane == 0) { L.score[k] = RA_SCORE_DROPPED; if (L.keep) L.keep[k] = 0; } continue; }   // nothing was evaluated; the entry point reports RA_EINVAL
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
	cudaError_t e = cudaMemsetAsync(P.totals, 0, 3 * sizeof(unsigned long long), s);   // parts, tiles, deferred parts
	if (e != cudaSuccess) return e;
	k_sp_prep<<<(L.n_groups + 255) / 256, 256, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

cudaError_t launch_sparse_eval(const ChainLaunch &L, const SparsePlan &P, int phase, int n_sm, cudaStream_t s)
{
	cudaError_t e = cudaMemsetAsync(L.work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	if (P.tile_parts == (uint32_t)kTileParts) {                              // CTA-wide tiles (test hook)
		if (L.cells) k_sp_parts<true><<<(uint32_t)n_sm * kSparseCtasPerSm, kSpThreads, 0, s>>>(L, P);
		else k_sp_parts<false><<<(uint32_t)n_sm * kSparseCtasPerSm, kSpThreads, 0, s>>>(L, P);
	} else {
		constexpr size_t kSmem = sizeof(Sp2Warp) * kSp2Warps;
		static unsigned long long attr_devices = 0;                          // > 48 KB of dynamic shared memory is opt-in, per function and device
		int dev = 0;
		cudaGetDevice(&dev);
		if (!((attr_devices >> (dev & 63)) & 1ull)) {
			e = cudaFuncSetAttribute(k_sp_parts2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
			if (e == cudaSuccess) e = cudaFuncSetAttribute(k_sp_parts2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
			if (e != cudaSuccess) return e;
			__atomic_fetch_or(&attr_devices, 1ull << (dev & 63), __ATOMIC_RELAXED);
		}
		if (L.cells) k_sp_parts2<true><<<(uint32_t)n_sm * kSparse2CtasPerSm, kSp2Warps * 32, kSmem, s>>>(L, P);
		else k_sp_parts2<false><<<(uint32_t)n_sm * kSparse2CtasPerSm, kSp2Warps * 32, kSmem, s>>>(L, P);
		if (P.rest) {
			if (L.cells) k_sp_rest<true><<<(uint32_t)n_sm * 8, 128, 0, s>>>(L, P);
			else k_sp_rest<false><<<(uint32_t)n_sm * 8, 128, 0, s>>>(L, P);
		}
	}
	k_sp_replay<<<(uint32_t)(((uint64_t)L.n_groups * 32 + 127) / 128), 128, 0, s>>>(L, P, phase);
	return cudaGetLastError();
}

} // namespace ra
