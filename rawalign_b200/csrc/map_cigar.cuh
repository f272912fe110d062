// map_cigar.cuh -- the --dtw-output-cigar stage of the map loop on the device (included by map.cu only).
// align_chain(cigar = true), rmap.cpp:218-235 / :282-299: the winning chain of every read that left the loop mapped is cut into
// the reference's sub-problems, every one runs DTW_global_tb (the generic traceback kernel, one warp per problem), the paths are
// compacted into one CSR array in chain order and shifted by the part's start anchor (:287-288; the global-border quirk of
// :230-233 -- only the LAST element is shifted, once per element -- is kept as a per-chain 64-bit addend), and the chain's
// alignment score is re-computed from the summed costs (:306).  Nothing but the finished paths and 12 bytes per chain goes back.
#pragma once

namespace mapk {

__global__ void k_cig_count(const ra_chain_t *__restrict__ ch, uint32_t n_fin, int sparse, uint32_t *__restrict__ parts)
{
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k < n_fin) parts[k] = sparse ? ch[k].n_anchors - 1u : 1u;
}

// one warp per chain, lanes over its parts (chain_items of api.cu on the device: part p runs from anchors[parts - p] to anchors[parts - p - 1])
__global__ void __launch_bounds__(128)
k_cig_items(const ra_chain_t *__restrict__ ch, const ra_anchor_t *__restrict__ anc, uint32_t n_fin, int sparse, const uint64_t *__restrict__ first,
            ra_dtw_item_t *__restrict__ items, uint32_t *__restrict__ slot_sz, uint32_t *__restrict__ dir_sz, uint32_t *__restrict__ maxw)
{
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (k >= n_fin) return;
	const ra_chain_t c = ch[k];
	const ra_anchor_t *a = anc + c.anchor_off;
	const uint32_t parts = sparse ? c.n_anchors - 1u : 1u;
	uint32_t mw = 0;
	for (uint32_t p = lane; p < parts; p += 32) {
		const ra_anchor_t s = sparse ? a[parts - p] : a[c.n_anchors - 1], e = sparse ? a[parts - p - 1] : a[0];
		ra_dtw_item_t it;
		it.a_off = c.read_off + s.query_position; it.b_off = c.ref_off + s.target_position;            // rmap.cpp:198-202, :256-260
		it.a_len = e.query_position - s.query_position + 1u; it.b_len = e.target_position - s.target_position + 1u;
		it.band_radius = -1; it.exclude_last = 0u;                                                      // DTW_global_tb; cigar ignores exclude_last (:283-284)
		const uint64_t q = first[k] + p;
		items[q] = it;
		const uint32_t diag = it.a_len + it.b_len - 1u;
		slot_sz[q] = diag;
		dir_sz[q] = (diag * ((it.b_len + 15u) >> 4) + 31u) & ~31u;
		mw = max(mw, it.b_len);
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) mw = max(mw, __shfl_xor_sync(0xffffffffu, mw, o));
	if (lane == 0 && mw) atomicMax(maxw, mw);
}

// one warp per chain: score (lane 0, the reference's float accumulation order) and the anchor shifts of its parts' paths
__global__ void __launch_bounds__(128)
k_cig_finish(const ra_chain_t *__restrict__ ch, const ra_anchor_t *__restrict__ anc, uint32_t n_fin, int sparse, const uint64_t *__restrict__ first,
             const ra_dtw_item_t *__restrict__ items, const float *__restrict__ cost, const uint64_t *__restrict__ path_off,
             uint32_t *__restrict__ out_i, uint32_t *__restrict__ out_j, float bonus, int contract,
             float *__restrict__ alns, uint64_t *__restrict__ chain_off, uint32_t *__restrict__ chain_len, uint64_t *__restrict__ last_add /* [2k], [2k+1] */)
{
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (k >= n_fin) return;
	const ra_chain_t c = ch[k];
	const ra_anchor_t *a = anc + c.anchor_off;
	const uint64_t f = first[k], l = first[k + 1];
	if (lane == 0) {
		float dtw_cost = 0.0f; uint32_t n_aligned = 0;
		for (uint64_t q = f; q < l; ++q) { dtw_cost = __fadd_rn(dtw_cost, cost[q]); n_aligned += items[q].a_len; }   // :226 / :290, :236 / :292
		alns[k] = contract ? __fmaf_rn((float)n_aligned, bonus, -dtw_cost) : __fsub_rn(__fmul_rn((float)n_aligned, bonus), dtw_cost);   // :306
		chain_off[k] = path_off[f]; chain_len[k] = (uint32_t)(path_off[l] - path_off[f]);
		uint64_t ai = 0, aj = 0;
		if (!sparse && l > f) {                                     // :230-233: alignment.back() is shifted once per path element
			const uint64_t len = path_off[l] - path_off[f];
			ai = len * (uint64_t)a[c.n_anchors - 1].query_position; aj = len * (uint64_t)a[c.n_anchors - 1].target_position;
		}
		last_add[2 * (size_t)k] = ai; last_add[2 * (size_t)k + 1] = aj;
	}
	if (!sparse) return;
	const uint32_t parts = c.n_anchors - 1u;
	for (uint64_t q = f; q < l; ++q) {
		const ra_anchor_t s = a[parts - (uint32_t)(q - f)];
		const uint64_t b = path_off[q], e = path_off[q + 1];
		for (uint64_t t = b + lane; t < e; t += 32) { out_i[t] += s.query_position; out_j[t] += s.target_position; }   // :287-288
	}
}

} // namespace mapk
