/* ra_oracle_map.h -- CPU restatement of the map loop around RawAlign's hot path.  TEST INFRASTRUCTURE ONLY. */
#ifndef RA_ORACLE_MAP_H
#define RA_ORACLE_MAP_H
#include <stdint.h>
#include "ra_oracle.h"
#ifdef __cplusplus
extern "C" {
#endif

#define ORA_M_DTW_EVALUATE_CHAINS 0x2   /* roptions.h:12-16 */
#define ORA_M_DTW_OUTPUT_CIGAR    0x4
#define ORA_M_DTW_LOG_SCORES      0x8
#define ORA_M_DISABLE_CHAININGSCORE_FILTERING 0x10

/* ri_mapopt_t, roptions.h:39-87, field for field */
typedef struct {
	uint32_t bp_per_sec, sample_rate, chunk_size;
	uint32_t min_events, max_gap_length, max_target_gap_length, chaining_band_length, max_num_skips, min_num_anchors, num_best_chains;
	float min_chaining_score;
	uint32_t step_size, max_num_chunk, min_chain_anchor, min_chain_anchor_out, dtw_border_constraint, dtw_fill_method;
	float dtw_band_radius_frac, dtw_match_bonus, dtw_min_score;
	float min_bestmap_ratio, min_bestmap_ratio_out, min_meanmap_ratio, min_meanmap_ratio_out;
	float t_threshold; uint32_t tn_samples, ttest_freq, tmin_reads;
	int64_t flag, mini_batch_size;
	uint32_t window_length1, window_length2; float threshold1, threshold2, peak_height;
} ora_mapopt_t;

/* the index, flattened: see ra_oracle_map.c */
typedef struct {
	int32_t w, e, n, q, lq, k;
	uint32_t n_seq;
	const uint32_t *seq_len;
	const float *const *fwd;     /* ri_idx_t::forward_signals (strand flag 1) */
	const float *const *rev;     /* ri_idx_t::reverse_signals (strand flag 0) */
	uint64_t n_keys;
	const uint64_t *keys, *key_off, *vals;
} ora_index_t;

/* ri_chain_t, rmap.h:30-46 */
typedef struct {
	float chaining_score, alignment_score;
	uint32_t reference_sequence_index, start_position, end_position, n_anchors;
	uint8_t mapq; int strand;
	ora_anchor_t *anchors;
} ora_chain_t;
typedef struct {
	float chaining_score, alignment_score; uint32_t reference_sequence_index, start_position, end_position, n_anchors;
	uint32_t mapq; int32_t strand; uint64_t anchor_off;
} ora_chain_out_t;

/* the numbers behind one PAF line (rmap.cpp:730-758, :954-972) */
typedef struct {
	uint32_t mapped, ref_id, read_length, read_start_position, read_end_position, fragment_start_position, fragment_length, mapq, rev;
	uint32_t ci, sl, cm, nc;
	float s1, s2, sm, at, aq, alns;
	uint32_t n_events; uint64_t cells;
} ora_paf_t;

struct ora_reader;
uint32_t ora_seq_to_sig(const uint8_t *seq, int len, const float *pore_vals, int k, int strand, int contract, float *out);
uint32_t ora_sketch(const float *ev, uint32_t len, int w, int e, int n, int q, int lq, int k, uint64_t *hash, uint32_t *pos, int *unsupported);
uint32_t ora_idx_get(const ora_index_t *ix, uint64_t h, const uint64_t **vals);
struct ora_reader *ora_reader_new(const ora_index_t *ix, const ora_mapopt_t *opt, int contract);
void ora_reader_free(struct ora_reader *R);
int ora_reader_chunk(struct ora_reader *R, const float *sig, uint32_t s_len);
int ora_reader_chunk_events(struct ora_reader *R, const float *cev, uint32_t n_ev);
uint32_t ora_reader_state(const struct ora_reader *R, uint32_t *n_events, uint32_t *offset, uint64_t *cells, uint64_t *n_seeds, uint64_t *n_hits);
const float *ora_reader_events(const struct ora_reader *R);
uint32_t ora_reader_chains(const struct ora_reader *R, int which, ora_chain_out_t *out, uint32_t cap, ora_anchor_t *anchors_out, uint64_t acap);
void ora_map_read(const ora_index_t *ix, const ora_mapopt_t *opt, int contract, const float *sig, uint32_t qlen, ora_paf_t *out);
void ora_paf_finish(const struct ora_reader *R, uint32_t qlen, uint32_t chunk_start, uint32_t current_chunk, ora_paf_t *o);
int ora_paf_format(const ora_paf_t *o, const char *read_name, const char *ref_name, uint32_t ref_len, char *buf, int cap);

#ifdef __cplusplus
}
#endif
#endif
