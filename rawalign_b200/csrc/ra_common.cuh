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

// Scratch layout of one sub-batch of the event pipeline (all strides in floats, multiples of 4).
struct EvScratch {
	uint32_t pstride, tstride, peak_stride;
	size_t floats_per_chunk;
};

// ---- launchers (events.cu) ---------------------------------------------------------------------
void events_scratch_layout(uint32_t max_len, const EvParams &p, EvScratch *s);
// Runs the five K1 kernels for n_chunks chunks.  scratch holds n_chunks * floats_per_chunk floats.
// ev_off: n_chunks+1 CSR offsets; with chained != 0 the scan continues from the value already in ev_off[0].
cudaError_t launch_events(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                          uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                          uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, cudaStream_t s,
                          cudaEvent_t after_detect);
cudaError_t launch_exclusive_scan_u32(const uint32_t *in, uint64_t *out, uint32_t n, cudaStream_t s);
cudaError_t launch_exclusive_scan_u32_chained(const uint32_t *in, uint64_t *out, uint32_t n, int chained, cudaStream_t s);
bool events_windows_supported(uint32_t w1, uint32_t w2);
// ---- launchers (events_fused.cu): one pass over the signal; windows up to kFusedMaxWindow ----------
constexpr uint32_t kFusedMaxWindow = 14;
constexpr int kFusedEventKernels = 3;
bool events_fused_supported(uint32_t w1, uint32_t w2);
void events_fused_layout(uint32_t max_len, const EvParams &p, uint32_t *raw_stride, size_t *floats_per_chunk);
cudaError_t launch_events_fused(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                                uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                                uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, int n_sm,
                                cudaStream_t s, cudaEvent_t after_detect);
constexpr int kEventKernels = 5;
// ---- launchers (events_pipe.cu): K1 v4, the default -- lane-per-chunk cp.async pipelines + parallel t-stat ------
constexpr uint32_t kPipeMaxWindow = 14;
constexpr int kPipeEventKernels = 5;
bool events_pipe_supported(uint32_t w1, uint32_t w2);
void events_pipe_layout(uint32_t max_len, const EvParams &p, uint32_t *stride, uint32_t *peak_stride, size_t *floats_per_chunk);
// optional second stream (created with a higher priority) for the two latency-bound kernels of the pipeline, and the four
// events that order them against the rest
struct PipeSplit { cudaStream_t hi; cudaEvent_t e[4]; };
extern cudaEvent_t *g_pipe_trace;             // RA_K1_TRACE: five timing events recorded by the next launch_events_pipe
cudaError_t launch_events_pipe(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                               uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                               uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, cudaStream_t s,
                               cudaEvent_t after_detect, cudaEvent_t scan_wait = nullptr, cudaEvent_t scan_done = nullptr, const PipeSplit *split = nullptr);

// ---- launcher (decode.cu): ri_read_sig's int16 -> pA conversion + (30,200) filter + compaction, one warp per read ----
cudaError_t launch_decode_raw(const int16_t *raw, const uint64_t *raw_off, const uint32_t *raw_len, const float *offset,
                              const float *scale, uint32_t n_reads, float *sig, uint32_t *l_sig, cudaStream_t s);

cudaError_t launch_fetch_tables(const uint64_t *off_host, const uint32_t *len_host, uint64_t *off_dev, uint32_t *len_dev, uint32_t n, cudaStream_t s);

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
bool dtw_tiny_supported(uint32_t a_len, uint32_t b_len, int32_t band_radius);      // R' <= 2: one item per thread
cudaError_t launch_dtw_tiny(const DtwLaunch &L, const uint32_t *order_dev, uint32_t n_order, cudaStream_t s);
cudaError_t launch_path_compact(const uint32_t *slot_i, const uint32_t *slot_j, const float *slot_d,
                                const uint64_t *slot_off, const ra_dtw_item_t *items, const uint32_t *path_len,
                                const uint64_t *path_off, uint32_t n_items, uint32_t *out_i, uint32_t *out_j,
                                float *out_d, cudaStream_t s);
uint32_t dtw_generic_smem_ring_cap();
// dtw_wave.cu: one warp per problem, register-resident skewed wavefront (H rows per lane; 0 = not eligible)
int dtw_wave_class(uint32_t a_len, uint32_t b_len, int32_t band_radius);
cudaError_t launch_dtw_wave(const DtwLaunch &L, int H, const uint32_t *order_dev, uint32_t n_order, uint32_t *work_counter, int n_sm, cudaStream_t s);

// ---- launcher (chains.cu) ----------------------------------------------------------------------
struct ChainLaunch {
	const float *read_arena, *ref_arena;
	const ra_anchor_t *anchors;
	const ra_chain_t *chains;
	const uint32_t *group_first;     // n_groups + 1 chain indices
	uint32_t n_groups;
	uint32_t border, fill, fp_contract;
	float band_frac, match_bonus, min_dtw_score;
	int min_score_mode;
	const float *min_score;          // per chain, mode 0; may be null
	float *score;
	uint8_t *keep;                   // may be null
	uint32_t *work_counter;
	float *ring_scratch;             // HBM rings for parts wider than the shared-memory ring; may be null
	uint32_t ring_cap;
	uint64_t *cells;                 // [issued, useful]; may be null
	uint32_t *need_cap;              // receives the ring capacity a skipped part needed (0 = none skipped)
	uint32_t *error;                 // set by k_validate_chains when a chain's anchors are not monotonic / out of the arenas; the
	                                 // evaluation kernels do nothing once it is set (the entry point returns RA_EINVAL)
	uint64_t n_read, n_ref;          // arena sizes in floats for the bounds check (0 = unknown: only monotonicity is checked)
};
// every anchor of every chain: anchors[0] is the chain end, positions do not increase along the array (rmap.cpp:456-465 makes
// them strictly monotonic), the chain end stays inside the arenas.  One warp per chain.
cudaError_t launch_validate_chains(const ChainLaunch &L, uint32_t k0, uint32_t k1, cudaStream_t s);
// global border through the wavefront item kernels (chains_global.cu)
cudaError_t launch_glob_prep(const ChainLaunch &L, int phase, const float *best, ra_dtw_item_t *items, int32_t *chain_item, uint32_t *n_items, cudaStream_t s);
cudaError_t launch_glob_replay(const ChainLaunch &L, int phase, float *best, const int32_t *chain_item, const float *cost, const unsigned long long *item_cells, cudaStream_t s);
constexpr int kChainCtasPerSm = 6;
cudaError_t launch_chain_eval(const ChainLaunch &L, int n_sm, cudaStream_t s);
// sparse border (chains_sparse.cu): flat part-cost kernel over tiles of 256 parts + replay kernel
struct SparsePlan {
	uint32_t n_chains;
	uint32_t *n_parts;               // per chain: parts evaluated in this phase (0 = chain inactive)
	uint64_t *part_base;             // per chain: first slot of its parts in part_cost / part_cells
	uint32_t *tile_chain, *tile_t0;  // per tile: chain and first part
	unsigned long long *totals;      // [0] parts, [1] tiles allocated so far in this phase, [2] parts deferred to k_sp_rest (device counters)
	float *part_cost;                // per evaluated part, at part_base[chain] + p
	uint32_t *part_cells;
	float *best;                     // per group: best score after phase A
	uint32_t tile_parts;             // parts per tile: kSparseTileParts2 (warp tiles, default) or kSparseTileParts (CTA tiles, test hook)
	uint64_t *rest;                  // chain << 32 | part of the parts deferred to k_sp_rest, counted in totals[2]; may be null
};
constexpr int kSparseCtasPerSm = 6;
#ifndef RA_SP2_CTAS
#define RA_SP2_CTAS 6                // with a staging window of 480 floats per span: 38 KB per CTA (profiles/r02_ab_sp2_stage_ctas.txt)
#endif
constexpr int kSparse2CtasPerSm = RA_SP2_CTAS;
constexpr int kSparseTileParts = 256;
constexpr int kSparseTileParts2 = 128;
constexpr int kSparseKernelsPerPhase = 4;   // prep, parts, rest, replay
// phase 0: min_score_mode 0; 1 / 2: first chain of every group / the remaining chains (min_score_mode 1)
cudaError_t launch_sparse_prep(const ChainLaunch &L, const SparsePlan &P, int phase, cudaStream_t s);
cudaError_t launch_sparse_eval(const ChainLaunch &L, const SparsePlan &P, int phase, int n_sm, cudaStream_t s);

} // namespace ra
