/*
 * ra_oracle.c -- CPU restatement of RawAlign's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the checker the CUDA path is compared with.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call it; the product
 * (rawalign_b200/, include/) never does.
 *
 * Parity status: PINNED.  Every function here is checked bit-for-bit against the reference's own
 * objects (dtw.cpp, revent.c compiled in place from /root/reference by oracle/Makefile into
 * oracle/_ref/) in tests/test_oracle_vs_ref.py, and against the committed fixtures generated from
 * that build (tests/golden/, made by oracle/gen_golden.py) on machines without /root/reference.
 *
 * Build with -ffp-contract=off (oracle/Makefile does): fused multiply-adds are written out
 * explicitly as fmaf()/fma() where the reference *build* has them, so the arithmetic does not
 * depend on what this compiler would contract.
 *
 * All file:line citations are relative to /root/reference/src.
 */
#include "ra_oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORA_INF 1e10f

/* ================================================================================================
 * Event detection -- revent.c
 * ============================================================================================== */

/* revent.c:22-32.  contract: the reference build fuses sig*sig + pss (revent.c:30). */
static void ora_prefix(const float *sig, uint32_t s_len, float *ps, float *pss, int contract)
{
	ps[0] = 0.0f;
	pss[0] = 0.0f;
	for (uint32_t i = 0; i < s_len; ++i) {
		ps[i + 1] = ps[i] + sig[i];
		if (contract) pss[i + 1] = fmaf(sig[i], sig[i], pss[i]);
		else { float sq = sig[i] * sig[i]; pss[i + 1] = pss[i] + sq; }
	}
}

/* revent.c:34-75.  tstat has s_len+1 entries, zero outside [w, s_len-w]. */
static void ora_tstat(const float *ps, const float *pss, uint32_t s_len, uint32_t w, float *tstat, int contract)
{
	memset(tstat, 0, (size_t)(s_len + 1) * sizeof(float));
	if (s_len < 2 * w || w < 2) return;                       /* :46 */
	const float wf = (float)w;
	for (uint32_t i = w; i <= s_len - w; ++i) {                /* :50 */
		float sum1 = ps[i], sumsq1 = pss[i];
		if (i > w) { sum1 -= ps[i - w]; sumsq1 -= pss[i - w]; } /* :53-56 */
		float sum2 = ps[i + w] - ps[i];
		float sumsq2 = pss[i + w] - pss[i];
		float mean1 = sum1 / wf, mean2 = sum2 / wf;            /* :59-60 */
		float var;
		float q1 = sumsq1 / wf, q2 = sumsq2 / wf;
		if (contract) {                                        /* :61, two fused negate-multiply-adds */
			float t1 = fmaf(-mean1, mean1, q1);
			float t2 = t1 + q2;
			var = fmaf(-mean2, mean2, t2);
		} else {
			float m1 = mean1 * mean1; float t1 = q1 - m1;
			float t2 = t1 + q2;
			float m2 = mean2 * mean2; var = t2 - m2;
		}
		var = fmaxf(var, FLT_MIN);                             /* :63 */
		const float delta_mean = mean2 - mean1;                /* :68 */
		const float vw = var / wf;
		tstat[i] = (float)(fabs((double)delta_mean) / sqrt((double)vw)); /* :69 (C: fabs/sqrt are double) */
	}
}

typedef struct {
	const float *sig; float threshold; uint32_t window_length; uint32_t masked_to;
	int peak_pos; float peak_value; int valid_peak;
} ora_detector;

/* revent.c:77-138 */
static uint32_t ora_gen_peaks(ora_detector *sd, ora_detector *ld, uint32_t s_len, float peak_height, uint32_t *peaks)
{
	uint32_t n = 0;
	ora_detector *det[2] = { sd, ld };
	for (uint32_t i = 0; i < s_len; i++) {
		for (int k = 0; k < 2; k++) {
			ora_detector *d = det[k];
			if (d->masked_to >= i) continue;                   /* :91 (unsigned compare) */
			float v = d->sig[i];
			if (d->peak_pos == -1) {                           /* :94 */
				if (v < d->peak_value) d->peak_value = v;
				else if (v - d->peak_value > peak_height) { d->peak_value = v; d->peak_pos = (int)i; }
			} else {
				if (v > d->peak_value) { d->peak_value = v; d->peak_pos = (int)i; }  /* :107-111 */
				if (d == sd && d->peak_value > d->threshold) { /* :113-120 */
					ld->masked_to = (uint32_t)d->peak_pos + d->window_length;
					ld->peak_pos = -1; ld->peak_value = FLT_MAX; ld->valid_peak = 0;
				}
				if (d->peak_value - v > peak_height && d->peak_value > d->threshold) d->valid_peak = 1; /* :122 */
				if (d->valid_peak && (i - (uint32_t)d->peak_pos) > d->window_length / 2) {              /* :126 */
					peaks[n++] = (uint32_t)d->peak_pos;
					d->peak_pos = -1; d->peak_value = v; d->valid_peak = 0;
				}
			}
		}
	}
	return n;
}

/* revent.c:140-188.  Returns n_ev; events written to ev. */
static uint32_t ora_gen_events(const uint32_t *peaks, uint32_t n_peaks, const float *ps, uint32_t s_len,
                               float *ev, int contract, double *mean_out, double *std_out)
{
	uint32_t n_ev = 1;
	double sum = 0, sum2 = 0;
	for (uint32_t i = 1; i < n_peaks; ++i)
		if (peaks[i] > 0 && peaks[i] < s_len) n_ev++;            /* :145-147 */
	float l_ps = 0, l_peak = 0;
	for (uint32_t pi = 0; pi + 1 < n_ev; pi++) {                 /* :160-167, peaks consumed in array order */
		float e = (ps[peaks[pi]] - l_ps) / ((float)peaks[pi] - l_peak);
		ev[pi] = e;
		sum += (double)e;
		float e2 = e * e;                                        /* float product, then widened (:164) */
		sum2 += (double)e2;
		l_ps = ps[peaks[pi]];
		l_peak = (float)peaks[pi];
	}
	{
		float e = (ps[s_len] - l_ps) / ((float)s_len - l_peak);  /* :172 */
		ev[n_ev - 1] = e;
		sum += (double)e;
		float e2 = e * e;
		sum2 += (double)e2;
	}
	double mean = sum / (double)n_ev;                            /* :179 */
	double q = sum2 / (double)n_ev;
	double sd;
	if (contract) sd = sqrt(fma(-mean, mean, q));                /* :180 fused in the reference build */
	else { double mm = mean * mean; sd = sqrt(q - mm); }
	for (uint32_t i = 0; i < n_ev; ++i)
		ev[i] = (float)(((double)ev[i] - mean) / sd);            /* :183 */
	if (mean_out) *mean_out = mean;
	if (std_out) *std_out = sd;
	return n_ev;
}

/* revent.c:190-210.  ev must hold s_len floats.  Returns the number of events (0 when the
 * reference returns NULL).  peaks_out (optional, s_len entries) receives the raw peak list. */
uint32_t ora_detect_events(const float *sig, uint32_t s_len, const ora_evopt_t *opt, int contract,
                           float *ev, uint32_t *peaks_out, uint32_t *n_peaks_out)
{
	if (s_len == 0) { if (n_peaks_out) *n_peaks_out = 0; return 0; }
	float *ps = (float *)calloc((size_t)s_len + 1, sizeof(float));
	float *pss = (float *)calloc((size_t)s_len + 1, sizeof(float));
	float *t1 = (float *)calloc((size_t)s_len + 1, sizeof(float));
	float *t2 = (float *)calloc((size_t)s_len + 1, sizeof(float));
	uint32_t *peaks = (uint32_t *)malloc((size_t)s_len * sizeof(uint32_t));
	ora_prefix(sig, s_len, ps, pss, contract);
	ora_tstat(ps, pss, s_len, opt->window_length1, t1, contract);
	ora_tstat(ps, pss, s_len, opt->window_length2, t2, contract);
	ora_detector sd = { t1, opt->threshold1, opt->window_length1, 0, -1, FLT_MAX, 0 };
	ora_detector ld = { t2, opt->threshold2, opt->window_length2, 0, -1, FLT_MAX, 0 };
	uint32_t n_peaks = ora_gen_peaks(&sd, &ld, s_len, opt->peak_height, peaks);
	uint32_t n_ev = 0;
	if (n_peaks > 0) n_ev = ora_gen_events(peaks, n_peaks, ps, s_len, ev, contract, 0, 0);
	if (peaks_out) memcpy(peaks_out, peaks, (size_t)n_peaks * sizeof(uint32_t));
	if (n_peaks_out) *n_peaks_out = n_peaks;
	free(ps); free(pss); free(t1); free(t2); free(peaks);
	return n_ev;
}

/* Batched convenience used by bench.py's cpu_baseline leg (CSR output like the C ABI). */
void ora_detect_events_batch(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len,
                             uint32_t n_chunks, const ora_evopt_t *opt, int contract,
                             float *ev_out, uint64_t *ev_off)
{
	uint64_t o = 0;
	for (uint32_t c = 0; c < n_chunks; ++c) {
		ev_off[c] = o;
		o += ora_detect_events(sig + chunk_off[c], chunk_len[c], opt, contract, ev_out + o, 0, 0);
	}
	ev_off[n_chunks] = o;
}

/* ================================================================================================
 * DTW -- dtw.cpp
 * ============================================================================================== */

static inline float ora_min3(float top, float left, float topleft)
{
	/* std::min(std::min(top,left),topleft): std::min(a,b) = (b<a) ? b : a   (dtw.cpp:52-55) */
	float m = (left < top) ? left : top;
	return (topleft < m) ? topleft : m;
}

/* DTW_global, dtw.cpp:37-66: rolling array over a, outer loop over b. */
float ora_dtw_global(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int exclude_last)
{
	float *dp = (float *)malloc((size_t)a_len * sizeof(float));
	dp[0] = fabsf(a[0] - b[0]);
	for (uint32_t j = 1; j < a_len; j++) dp[j] = dp[j - 1] + fabsf(a[j] - b[0]);
	for (uint32_t i = 1; i < b_len; i++) {
		float old_left = dp[0];
		dp[0] = dp[0] + fabsf(a[0] - b[i]);
		for (uint32_t j = 1; j < a_len; j++) {
			float top = dp[j - 1], left = dp[j], topleft = old_left;
			dp[j] = ora_min3(top, left, topleft) + fabsf(a[j] - b[i]);
			old_left = left;
		}
	}
	float res = dp[a_len - 1];
	free(dp);
	if (exclude_last) res -= fabsf(a[a_len - 1] - b[b_len - 1]);   /* :60-62 */
	return res;
}

/* Band geometry of DTW_global_slantedbanded_antidiagonalwise (dtw.cpp:284-302), after the swap
 * that makes n the longer length. */
void ora_band_geometry(uint32_t a_len, uint32_t b_len, int band_radius, uint32_t *n_out, uint32_t *m_out,
                       int *P_out, int *S_out)
{
	uint32_t n = a_len >= b_len ? a_len : b_len;
	uint32_t m = a_len >= b_len ? b_len : a_len;
	/* :298 -- evaluated in 32-bit unsigned arithmetic (uint32_t * int -> unsigned), then to int */
	int extra = (int)(((n - m) * (uint32_t)band_radius + n - 1u) / n);
	int R = band_radius + extra;
	*P_out = R + (R % 2 == 0 ? 1 : 0);
	*S_out = R + (R % 2 == 1 ? 1 : 0);
	*n_out = n; *m_out = m;
}

/* One anti-diagonal of the band as the closed interval [lo,hi] of row indices j (row = index into
 * the shorter sequence), already clipped to the matrix; lo > hi means empty. */
static inline void ora_diag_range(int start_i, int start_j, int len, int n, int m, int *lo, int *hi)
{
	/* cells (start_i - o, start_j + o), o in [0,len)  -> j in [start_j, start_j+len-1], i = start_i+start_j-j */
	int o0 = 0, o1 = len;                                     /* dtw.cpp:365-366 / :419-420 */
	if (start_i - (n - 1) > o0) o0 = start_i - (n - 1);
	if (-start_j > o0) o0 = -start_j;
	if (start_i + 1 < o1) o1 = start_i + 1;
	if (m - start_j < o1) o1 = m - start_j;
	*lo = start_j + o0; *hi = start_j + o1 - 1;
}

/* DTW_global_slantedbanded_antidiagonalwise, dtw.cpp:273-520, restated as a clean DP over the
 * cell set the reference visits: anti-diagonal d = i+j is visited exactly once, either as the
 * "primary" diagonal of column c (d = c + r(c), length P, centred on row r(c), :416-420) or as the
 * "secondary" one inserted when the centre row steps down (d = c + r(c) - 1, length S, :361-366);
 * r(c) advances when (r+1)*n <= m*c (:352-359).  Cells outside the set read as 1e10.
 * cells_out (optional) receives |cell set|. */
float ora_dtw_banded(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int band_radius,
                     int exclude_last, uint64_t *cells_out)
{
	if (a_len < b_len) { const float *t = a; a = b; b = t; uint32_t tl = a_len; a_len = b_len; b_len = tl; } /* :284-292 */
	uint32_t un, um; int P, S;
	ora_band_geometry(a_len, b_len, band_radius, &un, &um, &P, &S);
	const int n = (int)un, m = (int)um;
	/* three rolling anti-diagonals indexed by j+1 (slot 0 = row -1 = always outside) */
	float *buf = (float *)malloc((size_t)3 * (m + 2) * sizeof(float));
	float *d0 = buf, *d1 = buf + (m + 2), *d2 = buf + 2 * (m + 2);
	for (int k = 0; k < 3 * (m + 2); ++k) buf[k] = ORA_INF;
	int lo0 = 0, hi0 = -1, lo1 = 0, hi1 = -1, lo2 = 0, hi2 = -1; /* written ranges of d0 (cur), d1 (d-1), d2 (d-2) */
	uint64_t cells = 0;

	/* iteration 0 (:317-347): only the corner cell */
	d0[1] = fabsf(a[0] - b[0]); lo0 = hi0 = 0; cells = 1;

	int r = 0;
	for (int c = 1; c < n; ++c) {
		int inc = ((int64_t)(r + 1) * (int64_t)n <= (int64_t)m * (int64_t)c); /* :352-359 */
		if (inc) r++;
		for (int pass = inc ? 0 : 1; pass < 2; ++pass) {
			int start_i, start_j, len;
			if (pass == 0) { start_i = c + S / 2 - 1; start_j = r - S / 2; len = S; } /* :362-363 */
			else           { start_i = c + P / 2;     start_j = r - P / 2; len = P; } /* :416-417 */
			/* rotate: d2 <- d1 <- d0 <- (old d2, cleared) */
			float *t = d2; int tlo = lo2, thi = hi2;
			d2 = d1; lo2 = lo1; hi2 = hi1;
			d1 = d0; lo1 = lo0; hi1 = hi0;
			d0 = t;
			for (int j = tlo; j <= thi; ++j) d0[j + 1] = ORA_INF;
			ora_diag_range(start_i, start_j, len, n, m, &lo0, &hi0);
			const int d = start_i + start_j;
			for (int j = lo0; j <= hi0; ++j) {
				int i = d - j;
				/* neighbours (i, j-1) and (i-1, j) on diagonal d-1, (i-1, j-1) on d-2 */
				float v = ora_min3(d1[j], d1[j + 1], d2[j]) + fabsf(a[i] - b[j]);
				d0[j + 1] = v;
			}
			if (hi0 >= lo0) cells += (uint64_t)(hi0 - lo0 + 1);
		}
	}
	float res = d0[m];                                        /* cell (n-1, m-1), :506-512 */
	free(buf);
	if (cells_out) *cells_out = cells;
	if (exclude_last) res -= fabsf(a[n - 1] - b[m - 1]);       /* :514-516 */
	return res;
}

/* |cell set| only (no arithmetic): the GCUPS numerator. */
uint64_t ora_dtw_cells(uint32_t a_len, uint32_t b_len, int band_radius)
{
	if (band_radius < 0) return (uint64_t)a_len * b_len;
	uint32_t un, um; int P, S;
	ora_band_geometry(a_len, b_len, band_radius, &un, &um, &P, &S);
	const int n = (int)un, m = (int)um;
	uint64_t cells = 1; int r = 0, lo, hi;
	for (int c = 1; c < n; ++c) {
		int inc = ((int64_t)(r + 1) * (int64_t)n <= (int64_t)m * (int64_t)c);
		if (inc) {
			r++;
			ora_diag_range(c + S / 2 - 1, r - S / 2, S, n, m, &lo, &hi);
			if (hi >= lo) cells += (uint64_t)(hi - lo + 1);
		}
		ora_diag_range(c + P / 2, r - P / 2, P, n, m, &lo, &hi);
		if (hi >= lo) cells += (uint64_t)(hi - lo + 1);
	}
	return cells;
}

/* DTW_global_tb, dtw.cpp:595-667.  path_i/path_j/path_diff must hold a_len+b_len-1 entries;
 * returns the path length (forward order), cost in *cost_out. */
uint32_t ora_dtw_global_tb(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int exclude_last,
                           float *cost_out, uint32_t *path_i, uint32_t *path_j, float *path_diff)
{
	float *dp = (float *)malloc((size_t)a_len * b_len * sizeof(float));
#define DP(i, j) dp[(size_t)(i) * b_len + (j)]
	DP(0, 0) = fabsf(a[0] - b[0]);
	for (uint32_t i = 1; i < a_len; i++) DP(i, 0) = DP(i - 1, 0) + fabsf(a[i] - b[0]);
	for (uint32_t j = 1; j < b_len; j++) DP(0, j) = DP(0, j - 1) + fabsf(a[0] - b[j]);
	for (uint32_t i = 1; i < a_len; i++)
		for (uint32_t j = 1; j < b_len; j++)
			DP(i, j) = ora_min3(DP(i - 1, j), DP(i, j - 1), DP(i - 1, j - 1)) + fabsf(a[i] - b[j]);
	uint32_t i = a_len - 1, j = b_len - 1, len = 0;
	const uint32_t cap = a_len + b_len - 1;
	/* written back-to-front so the result is in forward order (:656-657) */
	uint32_t w = cap;
	--w; path_i[w] = i; path_j[w] = j; if (path_diff) path_diff[w] = fabsf(a[i] - b[j]); len++;
	while (i > 0 || j > 0) {                                   /* :624-654 */
		if (i == 0) j--;
		else if (j == 0) i--;
		else {
			float left = DP(i - 1, j), top = DP(i, j - 1), topleft = DP(i - 1, j - 1);
			float m_tt = (topleft < top) ? topleft : top;       /* min(top, topleft) */
			float m_lt = (topleft < left) ? topleft : left;     /* min(left, topleft) */
			if (left < m_tt) i--;
			else if (top < m_lt) j--;
			else { i--; j--; }
		}
		--w; path_i[w] = i; path_j[w] = j; if (path_diff) path_diff[w] = fabsf(a[i] - b[j]); len++;
	}
	float cost = DP(a_len - 1, b_len - 1);
#undef DP
	free(dp);
	if (w > 0) {
		memmove(path_i, path_i + w, (size_t)len * sizeof(uint32_t));
		memmove(path_j, path_j + w, (size_t)len * sizeof(uint32_t));
		if (path_diff) memmove(path_diff, path_diff + w, (size_t)len * sizeof(float));
	}
	if (exclude_last) { len--; cost -= fabsf(a[a_len - 1] - b[b_len - 1]); } /* :659-663 */
	*cost_out = cost;
	return len;
}

/* ================================================================================================
 * Chain evaluation -- rmap.cpp:181-313 (align_chain, cigar=false) and :515-524
 * ============================================================================================== */

/* align_chain with cigar=false.  anchors[0] is the chain END (rmap.cpp:194-196).  read/ref point at
 * the read's whole event array and the strand's whole reference signal.  score_fma: the reference
 * build may contract n_aligned*bonus - cost (rmap.cpp:306) into one fused multiply-subtract; pass 1
 * to mirror that.  cells_out (optional) accumulates the DP cells actually evaluated. */
float ora_align_chain(const ora_anchor_t *anchors, uint32_t n_anchors, const float *read, const float *ref,
                      const ora_dtwopt_t *opt, float min_score, int score_fma, uint64_t *cells_out)
{
	float dtw_cost = 0.0f;
	uint32_t n_aligned = 0;
	uint64_t cells = 0, cc;
	if (opt->dtw_border_constraint == 0) {                    /* GLOBAL, :192-237 */
		const ora_anchor_t s = anchors[n_anchors - 1], e = anchors[0];
		uint32_t ref_size = e.target_position - s.target_position + 1;
		uint32_t read_size = e.query_position - s.query_position + 1;
		float max_attainable = (float)read_size * opt->dtw_match_bonus; /* :205 */
		if (max_attainable < min_score) { if (cells_out) *cells_out += 0; return -1e10f; }
		if (opt->dtw_fill_method == 0) {
			dtw_cost = ora_dtw_global(read + s.query_position, read_size, ref + s.target_position, ref_size, 0);
			cells += (uint64_t)read_size * ref_size;
		} else {
			int R = (int)((float)read_size * opt->dtw_band_radius_frac); if (R < 1) R = 1; /* :214 */
			dtw_cost = ora_dtw_banded(read + s.query_position, read_size, ref + s.target_position, ref_size, R, 0, &cc);
			cells += cc;
		}
		n_aligned = read_size;
	} else {                                                  /* SPARSE, :238-300 */
		uint32_t parts = n_anchors - 1;
		const ora_anchor_t cs = anchors[n_anchors - 1], ce = anchors[0];
		uint32_t chain_read_size = ce.query_position - cs.query_position + 1;
		float cur_max = (float)chain_read_size * opt->dtw_match_bonus; /* :246 */
		for (uint32_t p = 0; p < parts; ++p) {
			const ora_anchor_t s = anchors[parts - p], e = anchors[parts - p - 1];
			uint32_t ref_size = e.target_position - s.target_position + 1;
			uint32_t read_size = e.query_position - s.query_position + 1;
			if (cur_max < min_score) { if (cells_out) *cells_out += cells; return -1e10f; } /* :265-268 */
			int excl = (p != parts - 1);
			float sub;
			if (opt->dtw_fill_method == 0) {
				sub = ora_dtw_global(read + s.query_position, read_size, ref + s.target_position, ref_size, excl);
				cells += (uint64_t)read_size * ref_size;
			} else {
				int R = (int)((float)read_size * opt->dtw_band_radius_frac); if (R < 1) R = 1; /* :276 */
				sub = ora_dtw_banded(read + s.query_position, read_size, ref + s.target_position, ref_size, R, excl, &cc);
				cells += cc;
			}
			dtw_cost += sub;                                  /* :279 */
			cur_max -= sub;                                   /* :280 */
			n_aligned += read_size;                           /* :292 */
		}
	}
	if (cells_out) *cells_out += cells;
	if (score_fma) return fmaf((float)n_aligned, opt->dtw_match_bonus, -dtw_cost);
	float prod = (float)n_aligned * opt->dtw_match_bonus;
	return prod - dtw_cost;                                   /* :306 */
}

/* rmap.cpp:515-524 for one read: chains already sorted by chaining score.  chain k has anchors
 * anchors[anchor_off[k] .. anchor_off[k+1]). */
void ora_evaluate_chains(const ora_anchor_t *anchors, const uint64_t *anchor_off, uint32_t n_chains,
                         const float *read, const float *const *ref_per_chain, const ora_dtwopt_t *opt,
                         int score_fma, float *alignment_score, uint8_t *keep, uint64_t *cells_out)
{
	float best = 0.0f;
	for (uint32_t k = 0; k < n_chains; ++k) {
		float s = ora_align_chain(anchors + anchor_off[k], (uint32_t)(anchor_off[k + 1] - anchor_off[k]),
		                          read, ref_per_chain[k], opt, best, score_fma, cells_out);
		alignment_score[k] = s;
		int kp = s >= opt->dtw_min_score;
		if (kp && s > best) best = s;
		if (keep) keep[k] = (uint8_t)kp;
	}
}

/* Batched DTW costs for bench.py's cpu_baseline leg: same item layout as the C ABI. */
void ora_dtw_batch(const float *a_arena, const float *b_arena, const ora_item_t *items, uint32_t n_items, float *cost)
{
	for (uint32_t k = 0; k < n_items; ++k) {
		const ora_item_t *it = &items[k];
		if (it->band_radius < 0)
			cost[k] = ora_dtw_global(a_arena + it->a_off, it->a_len, b_arena + it->b_off, it->b_len, (int)it->exclude_last);
		else
			cost[k] = ora_dtw_banded(a_arena + it->a_off, it->a_len, b_arena + it->b_off, it->b_len,
			                         it->band_radius, (int)it->exclude_last, 0);
	}
}

/* ================================================================================================
 * Checker for the kernel's division-by-window shortcut (rawalign_b200/csrc/events_fused.cu):
 *   rcp = RN(1/w); q0 = RN(x*rcp); r = fma(-q0, w, x); q = fma(r, rcp, q0)
 * must equal the correctly rounded x / w that revent.c:59-61,69 performs.  Exhaustive over all 2^23
 * mantissas at three exponents (the sequence is scale-invariant away from under/overflow; the kernel
 * only takes the shortcut for 1e-30 <= |x| <= 1e30).  Returns the number of mismatches.
 * ============================================================================================== */
uint64_t ora_check_fastdiv(uint32_t w)
{
	const float wf = (float)w;
	const float rcp = 1.0f / wf;
	uint64_t bad = 0;
	const float scales[3] = { 1.0f, 1048576.0f, 1.0f / 1048576.0f };
	for (int s = 0; s < 3; ++s)
		for (uint32_t mant = 0; mant < (1u << 23); ++mant) {
			union { uint32_t u; float f; } v;
			v.u = 0x3f800000u | mant;
			const float x = v.f * scales[s];
			const float q0 = x * rcp;
			const float r = fmaf(-q0, wf, x);
			const float q = fmaf(r, rcp, q0);
			if (q != x / wf) ++bad;
			const float qn = fmaf(fmaf(q0, wf, -x), rcp, -q0);   /* same sequence on -x */
			if (qn != (-x) / wf) ++bad;
		}
	return bad;
}

/* ==============================================================================================
 * Raw-signal decode: the conversion loop of ri_read_sig, src/rsig.cpp:213-222.
 *   float scale = s->ran/s->dig;                                         :215 (the caller passes it)
 *   float original_sig = (discretized_sig[i] + s->offset) * scale;      :217  two fp32 roundings
 *   if (original_sig > 30 && original_sig < 200) keep                    :218
 * discretized_sig is the int16 Signal dataset read into floats (exact).  Returns l_sig.
 * Parity note: ri_read_sig itself needs the HDF5 reader, which is not built here, so this restatement
 * is not pinned against a compiled reference object (stated in DESIGN.md).
 * ============================================================================================== */
uint32_t ora_decode_raw(const int16_t *raw, uint32_t n, float offset, float scale, float *out)
{
	uint32_t l_sig = 0;
	for (uint32_t i = 0; i < n; ++i) {
		const float d = (float)raw[i];
		const float t = d + offset;
		const float pa = t * scale;
		if (pa > 30 && pa < 200) out[l_sig++] = pa;
	}
	return l_sig;
}
