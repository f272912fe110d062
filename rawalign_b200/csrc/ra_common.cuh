// ra_common.cuh -- shared declarations of the sm_100a hot path (internal; the public surface is
// include/rawalign_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rawalign_b200.h"

namespace ra {

struct EvParams {
	uint32_t w1, w2;
	float thr1, thr2, peak_height;
};

// Per-chunk result of the scan kernel, consumed by the normalise/compact kernel.
struct EvChunkStat {
	double mean, stdv;
};

// ---- launchers (events.cu) ---------------------------------------------------------------------
// raw: [n_chunks][raw_stride] floats of un-normalised event means; n_ev: events per chunk.
cudaError_t launch_events_scan(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len,
                               uint32_t n_chunks, const EvParams &p, int contract, float *raw,
                               uint32_t raw_stride, uint32_t *n_ev, EvChunkStat *stat, cudaStream_t s);
cudaError_t launch_exclusive_scan_u32(const uint32_t *in, uint64_t *out, uint32_t n, cudaStream_t s);
cudaError_t launch_events_normalise(const float *raw, uint32_t raw_stride, const uint32_t *n_ev,
                                    const EvChunkStat *stat, const uint64_t *ev_off, uint32_t n_chunks,
                                    float *ev_out, uint64_t ev_cap, cudaStream_t s);
bool events_windows_supported(uint32_t w1, uint32_t w2);

// ---- launchers (dtw.cu) ------------------------------------------------------------------------
struct DtwLaunch {
	const float *a_arena, *b_arena;
	const ra_dtw_item_t *items;      // device
	const uint32_t *order;           // device, optional permutation (largest first); may be null
	uint32_t n_items;
	float *cost;                     // device
	uint32_t *work_counter;          // device, zeroed by the launcher
	float *ring_scratch;             // device scratch for items whose band does not fit shared memory
	uint32_t ring_cap;               // floats per ring (power of two) in ring_scratch, per warp
	// traceback only
	uint32_t *dirs;                  // 2-bit directions, per item at dir_off[k] (in 32-bit words)
	const uint64_t *dir_off;
	uint32_t *path_i, *path_j;       // per item slot of (a_len+b_len-1) entries at slot_off[k], filled from the end
	float *path_diff;
	const uint64_t *slot_off;
	uint32_t *path_len;
};
cudaError_t launch_dtw_generic(const DtwLaunch &L, bool traceback, int n_sm, cudaStream_t s, int *n_launches);
cudaError_t launch_path_compact(const uint32_t *slot_i, const uint32_t *slot_j, const float *slot_d,
                                const uint64_t *slot_off, const ra_dtw_item_t *items, const uint32_t *path_len,
                                const uint64_t *path_off, uint32_t n_items, uint32_t *out_i, uint32_t *out_j,
                                float *out_d, cudaStream_t s);
uint32_t dtw_generic_smem_ring_cap();

} // namespace ra
