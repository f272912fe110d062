// chains_global.cu -- global-border chain evaluation for device-resident chains (ra_align_chains_batch_dev, ra_map_batch):
// one DTW per chain over [first anchor .. last anchor] (src/rmap.cpp:192-237), evaluated by the register-wavefront item
// kernels (dtw_wave.cu) instead of the generic warp-per-chain kernel of chains.cu.
//
//   k_glob_prep    decides which chains are evaluated in this phase (the budget pre-check of :205-209 against the threshold
//                  known so far), and appends one ra_dtw_item_t per such chain (compacted with one atomic per warp);
//   (api.cu)       the items come to the host only to be sorted into wavefront classes (dtw_dev), the costs stay on the device;
//   k_glob_replay  one thread per group: the reference's serial float arithmetic in chain order -- pre-check, alignment_score =
//                  n_aligned * bonus - cost (:306, fused when the reference build fuses it), keep flag (:518), running
//                  best_found_alignment (:519-521).
// Phases as in chains_sparse.cu: 0 = explicit thresholds; 1 = first chain of every group (threshold 0.0f), 2 = the other
// chains whose budget is not already below the group's best after phase 1 (the running threshold only grows, so every chain
// the reference evaluates is evaluated; a few it would have dropped later are evaluated too and dropped by the replay).
#include "ra_common.cuh"

namespace ra {

__device__ __forceinline__ bool glob_active(const ChainLaunch &L, const float *best, int phase, uint32_t g, uint32_t k, uint32_t k0, float budget)
{
	if (phase == 1) return k == k0;
	if (phase == 0) return !(budget < (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED));
	return k != k0 && !(budget < best[g]);
}

__global__ void __launch_bounds__(256)
k_glob_prep(ChainLaunch L, int phase, const float *__restrict__ best, ra_dtw_item_t *__restrict__ items, int32_t *__restrict__ chain_item,
            uint32_t *__restrict__ n_items)
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	const int lane = threadIdx.x & 31;
	const bool invalid = *reinterpret_cast<volatile uint32_t *>(L.error) != 0;
	uint32_t k0 = 0, k1 = 0;
	if (g < L.n_groups && !invalid) { k0 = L.group_first[g]; k1 = L.group_first[g + 1]; }
	uint32_t mine = 0;
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *anc = L.anchors + ch.anchor_off;
		const uint32_t read_size = anc[0].query_position - anc[ch.n_anchors - 1].query_position + 1;       // :202
		mine += glob_active(L, best, phase, g, k, k0, (float)read_size * L.match_bonus) ? 1u : 0u;         // :205
	}
	uint32_t inc = mine;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
	uint32_t base = 0;
	if (lane == 31) base = atomicAdd(n_items, inc);
	base = __shfl_sync(0xffffffffu, base, 31) + inc - mine;
	for (uint32_t k = k0; k < k1; ++k) {
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *anc = L.anchors + ch.anchor_off;
		const ra_anchor_t s = anc[ch.n_anchors - 1], e = anc[0];                                            // :195-196
		const uint32_t read_size = e.query_position - s.query_position + 1;
		if (!glob_active(L, best, phase, g, k, k0, (float)read_size * L.match_bonus)) { if (phase != 2) chain_item[k] = -1; continue; }
		ra_dtw_item_t it;
		it.a_off = ch.read_off + s.query_position; it.b_off = ch.ref_off + s.target_position;              // :198-201
		it.a_len = read_size; it.b_len = e.target_position - s.target_position + 1;
		it.band_radius = (L.fill == RA_DTW_FILL_METHOD_BANDED) ? max(1, (int)((float)read_size * L.band_frac)) : -1;   // :214
		it.exclude_last = 0;
		items[base] = it; chain_item[k] = (int32_t)base; ++base;
	}
}

__global__ void __launch_bounds__(128)
k_glob_replay(ChainLaunch L, int phase, float *__restrict__ best_io, const int32_t *__restrict__ chain_item, const float *__restrict__ cost,
              const unsigned long long *__restrict__ item_cells)
{
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= L.n_groups) return;
	const bool invalid = *reinterpret_cast<volatile uint32_t *>(L.error) != 0;
	const uint32_t k0 = L.group_first[g], k1 = L.group_first[g + 1];
	const uint32_t kb = (phase == 2) ? k0 + 1 : k0, ke = (phase == 1) ? min(k0 + 1, k1) : k1;
	float best = (phase == 2) ? best_io[g] : 0.0f;                                                          // :515
	unsigned long long useful = 0;
	for (uint32_t k = kb; k < ke; ++k) {
		if (invalid) { L.score[k] = RA_SCORE_DROPPED; if (L.keep) L.keep[k] = 0; continue; }
		const ra_chain_t ch = L.chains[k];
		const ra_anchor_t *anc = L.anchors + ch.anchor_off;
		const uint32_t read_size = anc[0].query_position - anc[ch.n_anchors - 1].query_position + 1;
		const float ms = (L.min_score_mode == 1) ? best : (L.min_score ? L.min_score[k] : RA_SCORE_DROPPED);
		float score = RA_SCORE_DROPPED;
		if (!((float)read_size * L.match_bonus < ms)) {                                                     // :205-209
			const int32_t it = chain_item[k];
			const float c = cost[it];
			score = L.fp_contract ? __fmaf_rn((float)read_size, L.match_bonus, -c) : __fsub_rn(__fmul_rn((float)read_size, L.match_bonus), c);   // :236, :306
			if (item_cells) useful += item_cells[it];
		}
		const bool kp = score >= L.min_dtw_score;                                                           // :518
		L.score[k] = score; if (L.keep) L.keep[k] = kp ? 1 : 0;
		if (kp && score > best) best = score;                                                               // :519-521
	}
	if (phase == 1) best_io[g] = best;
	if (L.cells && useful) atomicAdd(reinterpret_cast<unsigned long long *>(L.cells + 1), useful);
}

cudaError_t launch_glob_prep(const ChainLaunch &L, int phase, const float *best, ra_dtw_item_t *items, int32_t *chain_item, uint32_t *n_items, cudaStream_t s)
{
	cudaError_t e = cudaMemsetAsync(n_items, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	k_glob_prep<<<(L.n_groups + 255) / 256, 256, 0, s>>>(L, phase, best, items, chain_item, n_items);
	return cudaGetLastError();
}

cudaError_t launch_glob_replay(const ChainLaunch &L, int phase, float *best, const int32_t *chain_item, const float *cost, const unsigned long long *item_cells, cudaStream_t s)
{
	k_glob_replay<<<(L.n_groups + 127) / 128, 128, 0, s>>>(L, phase, best, chain_item, cost, item_cells);
	return cudaGetLastError();
}

} // namespace ra
