/* ra_oracle.h -- CPU restatement of RawAlign's hot path.  TEST INFRASTRUCTURE ONLY (see ra_oracle.c). */
#ifndef RA_ORACLE_H
#define RA_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct { uint32_t window_length1, window_length2; float threshold1, threshold2, peak_height; } ora_evopt_t;
typedef struct { uint32_t dtw_border_constraint, dtw_fill_method; float dtw_band_radius_frac, dtw_match_bonus, dtw_min_score; } ora_dtwopt_t;
typedef struct { uint32_t target_position, query_position; } ora_anchor_t;
typedef struct { uint64_t a_off, b_off; uint32_t a_len, b_len; int32_t band_radius; uint32_t exclude_last; } ora_item_t;

uint32_t ora_detect_events(const float *sig, uint32_t s_len, const ora_evopt_t *opt, int contract,
                           float *ev, uint32_t *peaks_out, uint32_t *n_peaks_out);
void ora_detect_events_batch(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len,
                             uint32_t n_chunks, const ora_evopt_t *opt, int contract,
                             float *ev_out, uint64_t *ev_off);
float ora_dtw_global(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int exclude_last);
void ora_band_geometry(uint32_t a_len, uint32_t b_len, int band_radius, uint32_t *n_out, uint32_t *m_out, int *P_out, int *S_out);
float ora_dtw_banded(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int band_radius,
                     int exclude_last, uint64_t *cells_out);
uint64_t ora_dtw_cells(uint32_t a_len, uint32_t b_len, int band_radius);
uint32_t ora_dtw_global_tb(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int exclude_last,
                           float *cost_out, uint32_t *path_i, uint32_t *path_j, float *path_diff);
float ora_align_chain(const ora_anchor_t *anchors, uint32_t n_anchors, const float *read, const float *ref,
                      const ora_dtwopt_t *opt, float min_score, int score_fma, uint64_t *cells_out);
void ora_evaluate_chains(const ora_anchor_t *anchors, const uint64_t *anchor_off, uint32_t n_chains,
                         const float *read, const float *const *ref_per_chain, const ora_dtwopt_t *opt,
                         int score_fma, float *alignment_score, uint8_t *keep, uint64_t *cells_out);
void ora_dtw_batch(const float *a_arena, const float *b_arena, const ora_item_t *items, uint32_t n_items, float *cost);
uint64_t ora_check_fastdiv(uint32_t w);
uint32_t ora_decode_raw(const int16_t *raw, uint32_t n, float offset, float scale, float *out);

#ifdef __cplusplus
}
#endif
#endif
