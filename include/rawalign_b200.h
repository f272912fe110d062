/*
 * rawalign_b200.h -- C ABI of the B200 (sm_100a) hot path of RawAlign.
 *
 * The reference (CMU-SAFARI/RawAlign) has no plugin interface; the boundary is the two call
 * sites inside its per-read map loop:
 *
 *   float* detect_events(void *km, uint32_t s_len, const float* sig,
 *                        const ri_mapopt_t *opt, uint32_t *n);            src/revent.h:22, called rmap.cpp:551
 *   void   align_chain(ri_chain_t&, const ri_idx_t*, const float* read_events, uint32_t, uint32_t,
 *                      const ri_mapopt_t*, bool cigar, float min_score);   src/rmap.cpp:181, called :517, :716
 *       -> float      DTW_global(a, a_len, b, b_len, exclude_last)                              src/dtw.hpp:21
 *       -> float      DTW_global_slantedbanded_antidiagonalwise(a, a_len, b, b_len, R, excl)    src/dtw.hpp:25
 *       -> dtw_result DTW_global_tb(a, a_len, b, b_len, exclude_last)                           src/dtw.hpp:28
 *
 * Each entry point below is the *batched* form of one of those calls: a batch of signal chunks,
 * a batch of (read-event span, reference-signal span) pairs, or a batch of chains.  All pointers
 * are plain host pointers unless the name ends in _dev.  Every function returns 0 on success or a
 * negative RA_E* code; the message is available from ra_last_error().  The library never calls
 * exit() and never frees or retains caller memory (except the reference-signal mirror, which is
 * copied to the device in ra_ref_load).  There is no CPU fallback: without a CUDA device
 * ra_init fails with RA_ENODEV.
 */
#ifndef RAWALIGN_B200_H
#define RAWALIGN_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RA_OK        0
#define RA_ENODEV   -1   /* no CUDA device / wrong architecture */
#define RA_ECUDA    -2   /* a CUDA runtime call failed */
#define RA_EINVAL   -3   /* bad argument */
#define RA_ENOMEM   -4   /* host or device allocation failed */
#define RA_EUNSUP   -5   /* combination the reference itself aborts on (rmap.cpp:223-225, :301-304) */

/* same numeric values as roptions.h:20-26 */
#define RA_DTW_BORDER_CONSTRAINT_GLOBAL 0
#define RA_DTW_BORDER_CONSTRAINT_SPARSE 1
#define RA_DTW_BORDER_CONSTRAINT_LOCAL  2
#define RA_DTW_FILL_METHOD_FULL   0
#define RA_DTW_FILL_METHOD_BANDED 1

/* Sentinels of the reference: dtw.cpp uses 1e10 for "outside the band", rmap.cpp:207,266 uses
 * -1e10 for "chain dropped by early termination". */
#define RA_DTW_INF        1e10f
#define RA_SCORE_DROPPED -1e10f

/* The hot-path fields of ri_mapopt_t (roptions.h:39-87), same names, same defaults
 * (roptions.c:37-53).  fp_contract selects which build of revent.c is mirrored: 1 = the
 * reference Makefile's -O3 -march=native on an FMA host (GCC contracts revent.c:30, :61 twice and
 * :180 into fused multiply-adds), 0 = the same source built with -ffp-contract=off. */
typedef struct ra_mapopt_s {
	uint32_t window_length1;      /* 3 */
	uint32_t window_length2;      /* 6 */
	float    threshold1;          /* 4.30265f */
	float    threshold2;          /* 2.57058f */
	float    peak_height;         /* 1.0f */
	uint32_t dtw_border_constraint; /* RA_DTW_BORDER_CONSTRAINT_SPARSE */
	uint32_t dtw_fill_method;     /* RA_DTW_FILL_METHOD_BANDED */
	float    dtw_band_radius_frac;/* 0.10f */
	float    dtw_match_bonus;     /* 0.4f */
	float    dtw_min_score;       /* 20.0f */
	uint32_t fp_contract;         /* 1 */
	uint32_t reserved;            /* 0 = default kernels; test hooks (parity-tested alternatives): bit 0 = the earlier five-kernel
	                                 event detection (also the path taken for windows > 14), bit 1 = the single-pass fused
	                                 event kernel, bit 2 = every DTW item through the generic shared-memory kernel,
	                                 bit 3 = sparse chain parts through CTA-wide tiles instead of warp tiles */
} ra_mapopt_t;

/* One DTW problem = one call of DTW_global / DTW_global_slantedbanded_antidiagonalwise /
 * DTW_global_tb made by align_chain (rmap.cpp:211,215,221,273,277,284).  a = read-event span,
 * b = reference expected-signal span, both as float offsets into the two arenas passed with the
 * batch.  band_radius is the R align_chain computes (rmap.cpp:214,276) BEFORE the slant
 * extension of dtw.cpp:298-300; band_radius < 0 selects the full matrix (DTW_global). */
typedef struct ra_dtw_item_s {
	uint64_t a_off;
	uint64_t b_off;
	uint32_t a_len;
	uint32_t b_len;
	int32_t  band_radius;
	uint32_t exclude_last;
} ra_dtw_item_t;

/* ri_anchor_t, rmap.h:22-28 */
typedef struct ra_anchor_s {
	uint32_t target_position;
	uint32_t query_position;
} ra_anchor_t;

/* What align_chain reads from ri_chain_t (rmap.h:30-46) plus where the chain's read and
 * reference arrays live in the batch arenas.  Chains of one read must be contiguous and in the
 * order the reference evaluates them (descending chaining_score, rmap.cpp:512): the running
 * best_found_alignment threshold of rmap.cpp:515-524 is replayed per group. */
typedef struct ra_chain_s {
	uint64_t anchor_off;   /* first anchor of this chain in the anchors array (anchors[0] = chain END, rmap.cpp:194-196) */
	uint32_t n_anchors;
	uint32_t group;        /* chains with equal consecutive group ids share one best_found_alignment */
	uint64_t read_off;     /* float offset of the read's event array (p->events[read_id].values) in the read arena */
	uint64_t ref_off;      /* float offset of forward_signals[ref] / reverse_signals[ref] in the reference arena */
} ra_chain_t;

typedef struct ra_ctx ra_ctx;

/* ---- lifetime ------------------------------------------------------------------------------- */
int         ra_init(int device, ra_ctx **out);
void        ra_destroy(ra_ctx *ctx);
const char *ra_last_error(void);
/* ri_mapopt_init, roptions.c:5-61 (hot-path subset) */
void        ra_mapopt_init(ra_mapopt_t *opt);
int         ra_set_opt(ra_ctx *ctx, const ra_mapopt_t *opt);
int         ra_get_opt(const ra_ctx *ctx, ra_mapopt_t *opt);
/* Performance knobs that do not change results (defaults in parentheses): "chain_slice_min_anchors" (2097152: anchors per
 * upload slice of ra_align_chains_batch, at most 8 slices), "k1_groups" (8: chunk groups of the resident event detection that run
 * on concurrent streams), "k1_priority_stream" (1: the latency-bound K1 kernels on a high-priority stream), "map_sig_whole"
 * (0: the float form of ra_map_batch uploads per chunk round only the chunk each still-active read is at; 1: the whole batch
 * up front, which also leaves it resident for the device-resident form). */
int         ra_set_tuning(ra_ctx *ctx, const char *name, int64_t value);
/* number of kernels this context has launched since creation (bench.py's gpu_launches) */
uint64_t    ra_launch_count(const ra_ctx *ctx);
/* Device time of the most recent event-detection and DTW batches, measured with CUDA events on the
 * context's own stream: the detector kernel alone, all K1 kernels, all DTW kernels (milliseconds). */
int         ra_last_timing(ra_ctx *ctx, float *events_scan_ms, float *events_total_ms, float *dtw_ms);

/* ---- reference-signal mirror ---------------------------------------------------------------- */
/* Copies ri_idx_t::forward_signals / reverse_signals (rawindex.h:32-33) into one device arena:
 * [fwd(0) | rev(0) | fwd(1) | rev(1) ...].  arena_off receives 2*n_seq float offsets (fwd, rev per
 * sequence) to be used as ra_chain_t::ref_off / ra_dtw_item_t::b_off. */
int ra_ref_load(ra_ctx *ctx, uint32_t n_seq, const float *const *fwd, const float *const *rev,
                const uint32_t *len, uint64_t *arena_off);
/* Same, from one already concatenated host array (n_floats values). */
int ra_ref_load_flat(ra_ctx *ctx, const float *ref, uint64_t n_floats);
/* Device pointer / size of the mirror (for NCCL broadcast by the multi-GPU driver). */
int ra_ref_device(ra_ctx *ctx, float **dev_ptr, uint64_t *n_floats);
/* Allocate the mirror without filling it (non-root ranks before the broadcast). */
int ra_ref_alloc(ra_ctx *ctx, uint64_t n_floats);

/* ---- K1: event detection, batched detect_events (revent.c:190-210) ---------------------------- */
/* chunk c is sig[chunk_off[c] .. chunk_off[c]+chunk_len[c]).  Output is CSR: events of chunk c are
 * ev_out[ev_off[c] .. ev_off[c+1]); ev_off has n_chunks+1 entries.  A chunk in which the reference
 * finds no peak yields zero events (detect_events returns NULL and leaves *n at 0, revent.c:205-206).
 * ev_cap is the capacity of ev_out in floats; RA_EINVAL if it is too small (sum chunk_len is
 * always enough). */
int ra_detect_events_batch(ra_ctx *ctx, const float *sig, uint64_t n_sig,
                           const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                           float *ev_out, uint64_t ev_cap, uint64_t *ev_off);
/* Device-resident form: every pointer is a device pointer; runs on the context stream and
 * returns after the stream is synchronised. total_events (host) receives ev_off[n_chunks]. */
int ra_detect_events_batch_dev(ra_ctx *ctx, const float *sig_dev,
                               const uint64_t *chunk_off_dev, const uint32_t *chunk_len_dev,
                               uint32_t n_chunks, uint32_t max_chunk_len,
                               float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev,
                               uint64_t *total_events);
/* No-allocation form of ra_detect_events_batch_dev.  It has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and results (bit-identical) as ra_detect_events_batch_dev, every pointer a device pointer, but
 * the pipeline's scratch is a file-scope device array of 8 MiB that every GPU which loads the library holds until exit, and the
 * call makes no cudaMalloc / cudaFree or allocating library call.  It enqueues its kernels on `stream` (a cudaStream_t; NULL =
 * the default stream) and returns WITHOUT synchronising: ev_out_dev / ev_off_dev are valid once the stream reaches that point
 * (read ev_off_dev[n_chunks] for the event count).  Limit: n_chunks x scratch-per-chunk must fit the 8 MiB -- with the default
 * detector windows (3, 6) 100 chunks of up to 4096 samples, 399 of up to 1024; larger requests return RA_EINVAL with the
 * numbers in ra_last_error().  Windows above 14 return RA_EUNSUP.  Every call on a GPU shares the one array: calls that
 * overlap on the GPU (other streams, threads or contexts) race; serialise them. */
int ra_detect_events_noalloc(ra_ctx *ctx, const float *sig_dev, const uint64_t *chunk_off_dev, const uint32_t *chunk_len_dev,
                             uint32_t n_chunks, uint32_t max_chunk_len, float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev,
                             void *stream);

/* Raw-signal form: the conversion loop of ri_read_sig (src/rsig.cpp:213-222) and the chunk loop of
 * map_worker_for (src/rmap.cpp:684-689) moved to the device, so that the DAC samples cross the bus as int16.
 * Read r is raw[raw_off[r] .. raw_off[r]+raw_len[r]) (reads in order, not overlapping; 16-byte aligned starts are fastest);
 * offset[r] and scale[r] = range/digitisation are the float fields of ri_sig_t (rsig.h:16, rsig.cpp:215).
 * pA = ((float)raw + offset) * scale is kept iff 30 < pA < 200 (rsig.cpp:219-220); l_sig[r] receives the number of kept
 * samples (ri_sig_t::l_sig); the kept signal is cut into chunks of chunk_size samples, at most max_chunks per read.
 * chunk_first has n_reads+1 entries: chunks of read r are [chunk_first[r], chunk_first[r+1]); events come back in the same CSR
 * form as ra_detect_events_batch over those chunks (ev_off needs chunk_first[n_reads]+1 entries; sum over reads of
 * min(max_chunks, ceil(raw_len/chunk_size)) + 1 is always enough).  Copies are sliced and overlap the kernels. */
int ra_detect_events_raw_batch(ra_ctx *ctx, const int16_t *raw, uint64_t n_raw, const uint64_t *raw_off, const uint32_t *raw_len,
                               const float *offset, const float *scale, uint32_t n_reads, uint32_t chunk_size, uint32_t max_chunks,
                               uint32_t *l_sig, uint32_t *chunk_first, float *ev_out, uint64_t ev_cap, uint64_t *ev_off);

/* Extension (not in the reference's interface): the first sample, relative to its chunk, of every event
 * of the most recent event-detection call, in the same CSR order as ev_out.  Used by the synthetic
 * workload generator to place ground-truth anchors; the map loop does not need it. */
int ra_last_event_positions(ra_ctx *ctx, uint32_t *pos_out, uint64_t cap);

/* ---- K2/K3: DTW cost, batched DTW_global / DTW_global_slantedbanded_antidiagonalwise ---------- */
/* a_arena = read events, b_arena = reference signal; b_arena == NULL uses the mirror of
 * ra_ref_load.  cost_out[k] is bit-identical to the reference call for item k. */
int ra_dtw_batch(ra_ctx *ctx, const float *a_arena, uint64_t n_a, const float *b_arena, uint64_t n_b,
                 const ra_dtw_item_t *items, uint32_t n_items, float *cost_out);
int ra_dtw_batch_dev(ra_ctx *ctx, const float *a_arena_dev, const float *b_arena_dev,
                     const ra_dtw_item_t *items_dev, const ra_dtw_item_t *items_host,
                     uint32_t n_items, float *cost_out_dev);
/* Exact number of DP cells the reference evaluates for one item (the GCUPS numerator):
 * |cell set| of dtw.cpp:273-520 for band_radius >= 0, a_len*b_len otherwise. */
uint64_t ra_dtw_cells(uint32_t a_len, uint32_t b_len, int32_t band_radius);

/* ---- K4: DTW with traceback, batched DTW_global_tb (dtw.cpp:595-667) --------------------------- */
/* Always the full matrix (the reference has no banded traceback, rmap.cpp:223-225,284).
 * Path of item k is path_i/path_j[path_off[k] .. path_off[k+1]) in forward order, i indexing a
 * (read events) and j indexing b (reference), both relative to the item's spans; the third field
 * of alignment_element, |a[i]-b[j]|, is written to path_diff when it is non-NULL.
 * path_cap = capacity of the three path arrays; sum(a_len+b_len-1) is always enough. */
int ra_dtw_tb_batch(ra_ctx *ctx, const float *a_arena, uint64_t n_a, const float *b_arena, uint64_t n_b,
                    const ra_dtw_item_t *items, uint32_t n_items, float *cost_out,
                    uint32_t *path_i, uint32_t *path_j, float *path_diff, uint64_t path_cap,
                    uint64_t *path_off);

/* ---- chain level: batched align_chain + the evaluation loop of gen_chains ---------------------- */
/* For every chain: slices the sub-problems exactly as align_chain does for the border constraint
 * and fill method in the context options (rmap.cpp:192-300), evaluates all of them on the GPU in
 * one batch, then replays on the host, in the reference's order and float arithmetic, the early
 * termination against min_score (rmap.cpp:205-209, :265-268), the cost accumulation (:279-280) and
 * alignment_score = n_aligned*bonus - cost (:306).
 * min_score_mode 0: every chain is scored against the given min_score[k] (NULL = -1e10, the
 *                   default argument of align_chain);
 * min_score_mode 1: the running best_found_alignment of rmap.cpp:515-524 is replayed per group
 *                   (starts at 0.0f, raised by chains with score >= dtw_min_score).
 * alignment_score[k] receives chain.alignment_score; keep[k] (optional) receives
 * alignment_score >= dtw_min_score (rmap.cpp:518). cells_issued / cells_useful (optional) receive
 * the DP cells the GPU evaluated and the cells the reference would have evaluated.
 * Sparse border: a plan kernel, a flat part-cost kernel over all anchor-to-anchor parts and a replay kernel
 * (csrc/chains_sparse.cu); global border: one item per chain through the wavefront kernels with the replay on the host.
 * 8 bytes per anchor and 32 per chain go to the device, 5 bytes per chain come back.
 * Preconditions, checked (RA_EINVAL): chain records inside the anchor array; every chain's anchors monotonic (anchors[0] =
 * chain end, neither coordinate increases along the array -- the reference's chainer makes them strictly decreasing,
 * rmap.cpp:456-465); the chain end inside the read / reference arenas.  The interior anchors are checked on the device
 * (k_validate_chains) before any kernel indexes with them.  For the _dev form with caller-owned arenas the arena sizes are
 * unknown to the library: monotonicity is checked, the bounds of ra_chain_t::read_off / ref_off are the caller's. */
int ra_align_chains_batch(ra_ctx *ctx, const float *read_arena, uint64_t n_read,
                          const float *ref_arena, uint64_t n_ref,
                          const ra_anchor_t *anchors, uint64_t n_anchors,
                          const ra_chain_t *chains, uint32_t n_chains,
                          int min_score_mode, const float *min_score,
                          float *alignment_score, uint8_t *keep,
                          uint64_t *cells_issued, uint64_t *cells_useful);

/* read_arena == NULL uses the events the last ra_detect_events_batch left on the device (offsets in
 * ra_chain_t::read_off are then offsets into that call's ev_out): the map loop needs the events on the
 * host for seeding, the device keeps its copy for DTW.
 *
 * Device-resident form: all pointers are device pointers; group_first has n_groups+1 chain indices
 * (first chain of every group, then n_chains). read/ref NULL = resident events / reference mirror. */
int ra_align_chains_batch_dev(ra_ctx *ctx, const float *read_arena_dev, const float *ref_arena_dev,
                              const ra_anchor_t *anchors_dev, uint64_t n_anchors, const ra_chain_t *chains_dev, uint32_t n_chains,
                              const uint32_t *group_first_dev, uint32_t n_groups, int min_score_mode, const float *min_score_dev,
                              float *alignment_score_dev, uint8_t *keep_dev, uint64_t *cells_issued, uint64_t *cells_useful);

/* cigar=true form of align_chain (rmap.cpp:218-235, :282-299) for one batch of chains: the
 * concatenated alignment with anchor offsets applied exactly as the reference does (including
 * the quirks listed in SURVEY.md A.4: sparse ignores exclude_last_element and always uses the
 * full matrix; global offsets only the last element, n times).  RA_EUNSUP for global+banded. */
int ra_align_chains_cigar_batch(ra_ctx *ctx, const float *read_arena, uint64_t n_read,
                                const float *ref_arena, uint64_t n_ref,
                                const ra_anchor_t *anchors, uint64_t n_anchors,
                                const ra_chain_t *chains, uint32_t n_chains,
                                float *alignment_score, float *dtw_cost,
                                uint64_t *path_i, uint64_t *path_j, float *path_diff, uint64_t path_cap,
                                uint64_t *path_off);


/* ================================================================================================================
 * The map loop around the hot path on the device: seeding (ri_sketch, rsketch.c:223-274),
 * index lookup (ri_idx_get, rawindex.cpp:256-273), anchor sort and chaining DP (gen_chains, rmap.cpp:315-507),
 * chain evaluation (the hot path above), gen_primary_chains / comp_mapq (:65-128), the chunk loop with
 * is_mapped_with_high_confidence (:594-665, :685-693) and the PAF fields of map_worker_for (:696-800) -- for a whole batch
 * of reads, so that anchors and events never cross the bus: the int16 signal goes up, one ra_paf_t per read comes back.
 * ra_map_batch is the batched form of kt_for(n_threads, map_worker_for, step, n_sig) (rmap.cpp:916).
 * ================================================================================================================ */

/* ri_mapopt_t (roptions.h:39-87) field for field, so that a caller holding the reference's struct can pass it as is. */
typedef struct ra_ri_mapopt_s {
	uint32_t bp_per_sec, sample_rate, chunk_size;
	uint32_t min_events, max_gap_length, max_target_gap_length, chaining_band_length, max_num_skips, min_num_anchors, num_best_chains;
	float    min_chaining_score;
	uint32_t step_size, max_num_chunk, min_chain_anchor, min_chain_anchor_out, dtw_border_constraint, dtw_fill_method;
	float    dtw_band_radius_frac, dtw_match_bonus, dtw_min_score;
	float    min_bestmap_ratio, min_bestmap_ratio_out, min_meanmap_ratio, min_meanmap_ratio_out;
	float    t_threshold; uint32_t tn_samples, ttest_freq, tmin_reads;
	int64_t  flag, mini_batch_size;
	uint32_t window_length1, window_length2; float threshold1, threshold2, peak_height;
} ra_ri_mapopt_t;
#define RA_M_DTW_EVALUATE_CHAINS 0x2              /* roptions.h:12 */
#define RA_M_DTW_OUTPUT_CIGAR    0x4
#define RA_M_DISABLE_CHAININGSCORE_FILTERING 0x10
void ra_ri_mapopt_init(ra_ri_mapopt_t *opt);     /* ri_mapopt_init, roptions.c:5-61 */

/* The index on the device: an open-addressing table over the 32-bit seed hashes (built from the reference's unchanged
 * index file or from a flat table) plus the per-sequence signal offsets in the reference mirror, which it also loads. */
typedef struct ra_index ra_index;
/* Parses the file ri_idx_dump writes (rawindex.cpp:275-315): magic, (w,e,n,q,lq,k,n_seq,flag), per sequence name / len /
 * forward_signals / reverse_signals, then the 2^14 buckets.  Also loads the signals as the context's reference mirror. */
int ra_index_load_file(ra_ctx *ctx, const char *path, ra_index **out);
/* Same from memory: keys[n_keys] ascending unique seed hashes (riv.a[i].x >> RI_HASH_SHIFT), key_off[n_keys+1] into vals[],
 * vals = ref_id << 32 | position << 1 | strand (rsketch.h:40).  fwd/rev: ri_idx_t::forward_signals / reverse_signals. */
int ra_index_load_flat(ra_ctx *ctx, int32_t w, int32_t e, int32_t n, int32_t q, int32_t lq, int32_t k, uint32_t n_seq, const uint32_t *seq_len,
                       const float *const *fwd, const float *const *rev, uint64_t n_keys, const uint64_t *keys, const uint64_t *key_off,
                       const uint64_t *vals, ra_index **out);
/* Index construction on the device: what ri_idx_gen computes per sequence (rawindex.cpp:127-153:
 * ri_seq_to_sig for strand flag 1 then 0, rsig.cpp:7-41 -- fp64 mean / standard deviation summed in sequence order -- and
 * ri_sketch of both signals) and worker_post's grouping by seed (:194-247), as the device table ra_map_batch reads; the
 * expected signals become the context's reference mirror.  seq[i]: seq_len[i] base codes 0..3 = ACGT, 4 = ambiguous
 * (seq_nt4_table); pore_vals: 4^k level means (main.cpp:345-347).  The contraction mode follows the context's fp_contract.
 * Regular seeding only (w = n = 0).  ra_index_export returns the result as the flat table of ra_index_load_flat (positions in
 * ri_idx_get's order) -- bit-identical to the table parsed from an index file the reference built; keys == NULL sizes. */
int ra_index_build(ra_ctx *ctx, int32_t e, int32_t q, int32_t lq, int32_t k, uint32_t n_seq, const uint8_t *const *seq, const uint32_t *seq_len,
                   const float *pore_vals, ra_index **out);
int ra_index_export(ra_index *ix, uint64_t *n_keys, uint64_t *n_vals, uint64_t *keys, uint64_t *key_off, uint64_t *vals);
/* A copy of an index for another context -- a second batch in flight on the same GPU, or a peer GPU of the same process: the
 * three tables go device to device, the file is not parsed again. */
int ra_index_clone(ra_index *src, ra_ctx *ctx, ra_index **out);
void ra_index_destroy(ra_index *ix);
/* out[0..8] = w, e, n, q, lq, k, n_seq, n_keys (low 32 bits), n_vals (low 32 bits); seq_len_out (optional) n_seq entries */
int ra_index_info(const ra_index *ix, int32_t *out, uint32_t *seq_len_out);

/* Replication across the GPUs of a box (index and reference signals replicated, one broadcast at load): rank 0
 * loads the index; ra_index_tables gives the device pointers / byte sizes of its three tables {hash slots, positions, reference
 * mirror}; the other ranks call ra_index_alloc with rank 0's ra_index_info values, key / position counts and sequence lengths,
 * and receive the three tables into their own ra_index_tables pointers (ncclBroadcast; rawalign_b200/shard.py does it). */
int ra_index_tables(ra_index *ix, void **ptrs /* 3 */, uint64_t *bytes /* 3 */);
int ra_index_alloc(ra_ctx *ctx, const int32_t *info /* 9, as ra_index_info */, uint64_t n_keys, uint64_t n_vals, const uint32_t *seq_len, ra_index **out);

/* The numbers behind one PAF line (rmap.cpp:730-758, :954-972); the caller formats them (mt:f:, the per-read wall time, has no
 * per-read meaning in a batch). */
typedef struct ra_paf_s {
	uint32_t mapped, ref_id, read_length, read_start_position, read_end_position, fragment_start_position, fragment_length, mapq, rev;
	uint32_t ci, sl, cm, nc;          /* chunks used, l_sig, anchors of chain 0, chains */
	float    s1, s2, sm, at, aq, alns;
	uint32_t n_events;
	uint32_t reserved;
} ra_paf_t;

/* Per-stage device time of the most recent ra_map_batch (milliseconds, CUDA events) and its work counters. */
typedef struct ra_map_stats_s {
	float    decode_ms, events_ms, seed_ms, sort_ms, chain_ms, dtw_ms, select_ms, total_ms;
	uint32_t rounds; uint32_t reserved;
	uint64_t chunks, events, seeds, anchors, chains_evaluated, launches;
	uint64_t cells_issued, cells_useful;   /* DP cells of the chain evaluations; counted only when reserved == 1 on entry */
	uint64_t h2d_bytes, d2h_bytes;         /* bytes the call moved over PCIe (signal, tables, control words, records) */
} ra_map_stats_t;

/* Maps a batch of reads.  Signal either as int16 DAC samples (raw != NULL: read r is raw[raw_off[r] .. +raw_len[r]),
 * offset[r] / scale[r] as in ra_detect_events_raw_batch; decoded on the device, rsig.cpp:213-222) or as float pA
 * (raw == NULL: sig[sig_off[r] .. +sig_len[r]) = ri_sig_t::sig).  All pointers are host pointers.  out: n_reads records in
 * input order.  The float form moves, per chunk round, only the chunk each still-active read is at (sig_len[r] is l_sig, so
 * nothing else of a read that maps early is needed, rmap.cpp:685-712): read by the GPU straight from sig when that is
 * page-locked memory, else gathered into a page-locked staging buffer first; sig must stay valid and unchanged during the call.
 * raw == NULL and sig == NULL maps again the batch the previous raw-form (or "map_sig_whole") call left on the device (no
 * upload, no decode: the device-resident form).  Minimizer (w > 0) and BLEND (n > 0) indexes return RA_EUNSUP. */
int ra_map_batch(ra_ctx *ctx, ra_index *ix, const ra_ri_mapopt_t *opt,
                 const int16_t *raw, const uint64_t *raw_off, const uint32_t *raw_len, const float *offset, const float *scale,
                 const float *sig, const uint64_t *sig_off, const uint32_t *sig_len,
                 uint32_t n_reads, ra_paf_t *out, ra_map_stats_t *stats);

/* With RA_M_DTW_OUTPUT_CIGAR in opt->flag, ra_map_batch also runs align_chain(chains[0], cigar=true) for every mapped read
 * (rmap.cpp:715-717; full matrix + traceback per anchor-to-anchor part, the quirks of SURVEY.md A.4 included) and
 * ra_paf_t::alns is the score that call leaves in chains[0].alignment_score (what the alns:f: tag prints, :741).  The
 * alignments of the last such call, CSR by read: elements [path_off[r], path_off[r+1]) of path_i (read event), path_j
 * (reference position), path_diff -- the triples of the aln:s: tag (:580-592).  Call with path_i == NULL to size (path_off is
 * filled, n_reads+1 entries). */
int ra_map_last_cigar(ra_index *ix, uint32_t n_reads, uint64_t *path_off, uint64_t *path_i, uint64_t *path_j, float *path_diff, uint64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* RAWALIGN_B200_H */
