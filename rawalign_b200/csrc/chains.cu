// chains.cu -- batched align_chain(cigar=false) + the evaluation loop of gen_chains, entirely on the
// device (reference: src/rmap.cpp:181-313 and :515-524).
//
// One warp owns one *group* (the chains of one read in one chunk round, in the reference's
// evaluation order) and walks its chains sequentially, because chain k's early-termination
// threshold is the best score of chains 0..k-1 (rmap.cpp:517-521).  Within a chain:
//   * the sub-problems ("parts": anchor-to-anchor spans in sparse mode, one region in global mode)
//     are derived from the anchors on the fly -- 8 bytes per part cross PCIe, not a 32-byte descriptor;
//   * small parts (widest anti-diagonal <= 8 cells: every anchor gap below ~80 events at banded=0.1)
//     are solved one per lane, three rolling anti-diagonals per lane in shared memory;
//   * larger parts are solved by the whole warp (lanes stride over the anti-diagonal);
//   * parts are processed in tiles of kTile; after each tile lane 0 replays the reference's sequential
//     float arithmetic (cost += sub, budget -= sub, rmap.cpp:279-280) and its early exit (:265-268), so
//     a dropped chain stops issuing work at the next tile, as the reference stops at the next part.
// Output: alignment_score and keep flag per chain; nothing else returns to the host.
#include "dtw_core.cuh"

namespace ra {

constexpr int kChainWarps = 4;
constexpr int kTile = 256;                       // parts per replay tile
constexpr uint32_t kWarpRingCap = 256;           // floats per ring for warp-cooperative parts in shared memory

struct PartDesc { const float *A, *B; int n, m, band; uint32_t a_len; bool excl; };

__device__ __forceinline__ PartDesc make_part(const ChainLaunch &L, const ra_chain_t &ch, const ra_anchor_t *anc, uint32_t parts, uint32_t p)
{
	// sparse: part p runs from anchors[parts-p] to anchors[parts-p-1]  (rmap.cpp:253-260)
	// global: the single part runs from anchors[n-1] to anchors[0]     (rmap.cpp:195-202)
	ra_anchor_t s, e;
	if (L.border == RA_DTW_BORDER_CONSTRAINT_GLOBAL) { s = anc[ch.n_anchors - 1]; e = anc[0]; }
	else { s = anc[parts - p]; e = anc[parts - p - 1]; }
	PartDesc d;
	const uint32_t a_len = e.query_position - s.query_position + 1;
	const uint32_t b_len = e.target_position - s.target_position + 1;
	d.a_len = a_len;
	const float *a = L.read_arena + ch.read_off + s.query_position;
	const float *b = L.ref_arena + ch.ref_off + s.target_position;
	d.band = (L.fill == RA_DTW_FILL_METHOD_BANDED) ? max(1, (int)((float)a_len * L.band_frac)) : -1;   // :214, :276
	d.excl = (L.border == RA_DTW_BORDER_CONSTRAINT_SPARSE) && (p != parts - 1);                         // :270
	if (a_len >= b_len) { d.A = a; d.n = (int)a_len; d.B = b; d.m = (int)b_len; }                        // dtw.cpp:284-292
	else { d.A = b; d.n = (int)b_len; d.B = a; d.m = (int)a_len; }
	return d;
}

__global__ void __launch_bounds__(kChainWarps * 32)
k_chain_eval(ChainLaunch L)
{
	__shared__ float lane_ring[kChainWarps][2 * kNarrowRing * 32];   // [row ring | B ring], lane-interleaved
	__shared__ float warp_ring[kChainWarps][3 * kWarpRingCap];
	__shared__ float tile_cost[kChainWarps][kTile];
	__shared__ uint32_t tile_alen[kChainWarps][kTile];
	__shared__ uint32_t tile_cells[kChainWarps][kTile];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const uint32_t gwarp = blockIdx.x * kChainWarps + wid;
	float *my_lane_ring = &lane_ring[wid][lane];
	float *my_gring = L.ring_scratch ? L.ring_scratch + (size_t)gwarp * 3 * L.ring_cap : nullptr;
	uint64_t cells_issued = 0, cells_useful = 0;

	if (*reinterpret_cast<volatile uint32_t *>(L.error)) return;             // a chain failed validation: nothing is evaluated
	for (;;) {
		uint32_t g = 0;
		if (lane == 0) g = atomicAdd(L.work_counter, 1u);
		g = __shfl_sync(0xffffffffu, g, 0);
		if (g >= L.n_groups) break;
		const uint32_t k0 = L.group_first[g], k1 = L.group_first[g + 1];
		float best = 0.0f;                                                   // rmap.cpp:515
		for (uint32_t k = k0; k < k1; ++k) {
			const ra_chain_t ch = L.chains[k];
			const ra_anchor_t *anc = L.anchors + ch.anchor_off;
			const float ms = (L.min_score_mode == 1) ? best : (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED);
			const bool global = (L.border == RA_DTW_BORDER_CONSTRAINT_GLOBAL);
			const uint32_t parts = global ? 1u : ch.n_anchors - 1u;
			const ra_anchor_t a_first = anc[ch.n_anchors - 1], a_last = anc[0];
			const uint32_t chain_read_size = a_last.query_position - a_first.query_position + 1;   // :202, :245
			float cur_max = (float)chain_read_size * L.match_bonus;          // :205, :246
			float dtw_cost = 0.0f;
			uint32_t n_aligned = 0;
			bool dropped = false;
			if (global && cur_max < ms) dropped = true;                       // :206-209 (sparse checks inside the part loop)
			for (uint32_t t0 = 0; t0 < parts && !dropped; t0 += kTile) {
				const uint32_t tn = min((uint32_t)kTile, parts - t0);
				if (!global && cur_max < ms) { dropped = true; break; }       // the check before part t0 (:265-268)
				// ---- small parts: one per lane ----
				uint32_t big_mask_any = 0;
				for (uint32_t base = 0; base < tn; base += 32) {
					const uint32_t p = t0 + base + lane;
					bool is_big = false;
					if (base + lane < tn) {
						const PartDesc d = make_part(L, ch, anc, parts, p);
						tile_alen[wid][base + lane] = d.a_len;
						if (narrow_ok(d.n, d.m, d.band)) {
							uint32_t cnt = 0;
							float c = thread_dtw_narrow(LdgLoader{d.A}, d.n, LdgLoader{d.B}, d.m, d.band, my_lane_ring, my_lane_ring + kNarrowRing * 32, &cnt);
							if (d.excl) c -= fabsf(__ldg(d.A + d.n - 1) - __ldg(d.B + d.m - 1));   // dtw.cpp:514-516 / :60-62
							tile_cost[wid][base + lane] = c;
							tile_cells[wid][base + lane] = cnt;
							cells_issued += cnt;
						} else is_big = true;
					}
					// ---- large parts of this 32-block: the whole warp, one part at a time ----
					uint32_t bm = __ballot_sync(0xffffffffu, is_big);
					big_mask_any |= bm;
					while (bm) {
						const int src = __ffs(bm) - 1; bm &= bm - 1;
						const PartDesc d = make_part(L, ch, anc, parts, t0 + base + src);
						const int width = band_width_of(d.n, d.m, d.band);
						uint32_t cap = 1; while ((int)cap < width) cap <<= 1;
						float *ring = warp_ring[wid];
						if (cap > kWarpRingCap) {
							if (!my_gring || cap > L.ring_cap) {          // the host retries with a larger HBM ring
								if (lane == 0) { atomicMax(L.need_cap, cap); tile_cost[wid][base + src] = 0.0f; tile_cells[wid][base + src] = 0; }
								continue;
							}
							ring = my_gring; cap = L.ring_cap;
						} else cap = kWarpRingCap;
						uint64_t cnt = 0;
						float c = warp_dtw_cost(d.A, d.n, d.B, d.m, d.band, ring, cap, lane, &cnt);
						if (d.excl) c -= fabsf(__ldg(d.A + d.n - 1) - __ldg(d.B + d.m - 1));
						if (lane == 0) { tile_cost[wid][base + src] = c; tile_cells[wid][base + src] = (uint32_t)min(cnt, (uint64_t)0xffffffffu); cells_issued += cnt; }
					}
				}
				__syncwarp();
				// ---- replay, lane 0, reference order and arithmetic ----
				if (lane == 0) {
					if (global) {
						dtw_cost = tile_cost[wid][0]; n_aligned = tile_alen[wid][0];               // :211-216, :236
						cells_useful += tile_cells[wid][0];
					} else {
						for (uint32_t q = 0; q < tn; ++q) {
							if (cur_max < ms) { dropped = true; break; }                            // :265-268
							const float sub = tile_cost[wid][q];
							dtw_cost += sub;                                                        // :279
							cur_max -= sub;                                                         // :280
							n_aligned += tile_alen[wid][q];                                         // :292
							cells_useful += tile_cells[wid][q];
						}
					}
				}
				dropped = __shfl_sync(0xffffffffu, (int)dropped, 0) != 0;
				cur_max = __shfl_sync(0xffffffffu, cur_max, 0);
				__syncwarp();
			}
			float score = RA_SCORE_DROPPED;
			if (lane == 0) {
				if (!dropped) {
					// rmap.cpp:306; the reference build fuses the multiply and the subtract when fp_contract is set
					score = L.fp_contract ? __fmaf_rn((float)n_aligned, L.match_bonus, -dtw_cost)
					                      : __fsub_rn(__fmul_rn((float)n_aligned, L.match_bonus), dtw_cost);
				}
				const bool kp = score >= L.min_dtw_score;                                           // :518
				L.score[k] = score;
				if (L.keep) L.keep[k] = kp ? 1 : 0;
				if (kp && score > best) best = score;                                               // :519-521
			}
			best = __shfl_sync(0xffffffffu, best, 0);
		}
	}
	// cell counters (issued = evaluated on the GPU); lane-private counts are reduced per warp
	if (L.cells) {
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) cells_issued += __shfl_down_sync(0xffffffffu, cells_issued, o);
		if (lane == 0) {
			atomicAdd(reinterpret_cast<unsigned long long *>(L.cells), (unsigned long long)cells_issued);
			atomicAdd(reinterpret_cast<unsigned long long *>(L.cells + 1), (unsigned long long)cells_useful);
		}
	}
}

__global__ void __launch_bounds__(128)
k_validate_chains(ChainLaunch L, uint32_t k0, uint32_t k1)
{
	const uint32_t k = k0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	const int lane = threadIdx.x & 31;
	if (k >= k1) return;
	const ra_chain_t ch = L.chains[k];
	const ra_anchor_t *__restrict__ anc = L.anchors + ch.anchor_off;
	bool bad = ch.n_anchors == 0;
	for (uint32_t p = 1 + lane; p < ch.n_anchors; p += 32) {
		const ra_anchor_t a = anc[p - 1], b = anc[p];
		bad |= (b.query_position > a.query_position) | (b.target_position > a.target_position);
	}
	if (lane == 0 && ch.n_anchors) {
		const ra_anchor_t e = anc[0];
		if (L.n_read && ch.read_off + e.query_position >= L.n_read) bad = true;
		if (L.n_ref && ch.ref_off + e.target_position >= L.n_ref) bad = true;
	}
	if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(L.error, 1u);
}

cudaError_t launch_validate_chains(const ChainLaunch &L, uint32_t k0, uint32_t k1, cudaStream_t s)
{
	if (k1 <= k0) return cudaSuccess;
	k_validate_chains<<<(uint32_t)(((uint64_t)(k1 - k0) * 32 + 127) / 128), 128, 0, s>>>(L, k0, k1);
	return cudaGetLastError();
}

cudaError_t launch_chain_eval(const ChainLaunch &L, int n_sm, cudaStream_t s)
{
	if (L.n_groups == 0) return cudaSuccess;
	cudaError_t e = cudaMemsetAsync(L.work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	if (L.cells) { e = cudaMemsetAsync(L.cells, 0, 2 * sizeof(uint64_t), s); if (e != cudaSuccess) return e; }
	e = cudaMemsetAsync(L.need_cap, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	uint32_t ctas = (uint32_t)n_sm * kChainCtasPerSm;
	const uint32_t need = (L.n_groups + kChainWarps - 1) / kChainWarps;
	if (ctas > need) ctas = need;
	k_chain_eval<<<ctas, kChainWarps * 32, 0, s>>>(L);
	return cudaGetLastError();
}

} // namespace ra
