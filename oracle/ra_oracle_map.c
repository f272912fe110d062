/*
 * ra_oracle_map.c -- CPU restatement of the map loop around RawAlign's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Rows f1 (chaining), f2 (sketch + index lookup), f4 (reference -> expected signal) of SURVEY.md section 8(f) and the
 * per-read control flow that ties them to the hot path (gen_chains, gen_primary_chains, comp_mapq,
 * is_mapped_with_high_confidence, map_worker_for).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may
 * build, load or call it.
 *
 * Parity status: PINNED.  tests/test_map_oracle.py diffs every stage against the reference's own compiled code
 * (oracle/_ref/libra_refmap.so: ri_sketch, ri_idx_get, ri_seq_to_sig, gen_chains through ri_map_frag, map_worker_for) where
 * that is built, and against the committed fixtures generated from it (tests/golden/map_cases.npz) elsewhere.
 *
 * Scope limits (stated, not silent): minimizer sketching (w > 0, rsketch.c:131-221, the "faster" preset) and the BLEND
 * mode (n > 0, disabled in the reference itself) are not restated; ora_sketch returns 0 seeds with *unsupported = 1.
 *
 * All file:line citations are relative to /root/reference/src.
 */
#include "ra_oracle.h"
#include "ra_oracle_map.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ================================================================================================
 * f4: reference sequence -> expected signal, rsig.cpp:7-41
 * ============================================================================================== */
/* seq: 2-bit codes 0..3 (A,C,G,T), 4 = ambiguous (seq_nt4_table).  strand 1 walks the sequence backwards with
 * complemented bases (rsig.cpp:15-22).  fp64 sum / sum of squares in sequence order; contract: the reference build
 * fuses curval*curval + sum2 (:31) and sum2/j - mean*mean (:35).  Returns s_len = len - k + 1. */
uint32_t ora_seq_to_sig(const uint8_t *seq, int len, const float *pore_vals, int k, int strand, int contract, float *out)
{
	const uint64_t mask = (1ULL << (2 * k)) - 1;
	uint64_t kmer = 0;
	double sum = 0, sum2 = 0;
	int j = 0;
	for (int i = 0; i < len; ++i) {
		int pos = strand ? len - i - 1 : i;
		int c = seq[pos];
		if (c < 4) kmer = strand ? (((kmer << 2) | (3ULL ^ (uint64_t)c)) & mask) : (((kmer << 2) | (uint64_t)c) & mask);
		else kmer = (kmer << 2) & mask;
		if (i + 1 < k) continue;
		double curval = pore_vals[kmer];
		out[j++] = (float)curval;
		sum += curval;
		if (contract) sum2 = fma(curval, curval, sum2); else sum2 += curval * curval;
	}
	double mean = sum / j;
	double q = sum2 / j;
	double var = contract ? fma(-mean, mean, q) : q - mean * mean;
	double sd = sqrt(var);
	for (int i = 0; i < j; ++i) out[i] = (float)(((double)out[i] - mean) / sd);
	return (uint32_t)j;
}

/* ================================================================================================
 * f2: sketch (rsketch.c:5-15, :223-283) and index lookup (rawindex.cpp:256-273)
 * ============================================================================================== */
static inline uint64_t ora_hash64(uint64_t key, uint64_t mask)          /* rsketch.c:6-15 */
{
	key = (~key + (key << 21)) & mask;
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8)) & mask;
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4)) & mask;
	key = key ^ key >> 28;
	key = (key + (key << 31)) & mask;
	return key;
}

/* ri_sketch_reg, rsketch.c:223-274, with id = 0, strand = 0 as rmap.cpp:367 calls it.  Seed s: hash[s] = x >> RI_HASH_SHIFT
 * (what rmap.cpp:372 looks up), pos[s] = (uint32_t)y >> RI_POS_SHIFT (:373) = index of the e-mer's LAST event. */
uint32_t ora_sketch(const float *ev, uint32_t len, int w, int e, int n, int q, int lq, int k, uint64_t *hash, uint32_t *pos,
                    int *unsupported)
{
	(void)k;
	if (unsupported) *unsupported = 0;
	if (w != 0 || n != 0) { if (unsupported) *unsupported = 1; return 0; }
	const uint32_t quant_bit = (uint32_t)lq + 2, shift_r = 32 - (uint32_t)q;
	const uint64_t mask = (1ULL << 32) - 1, mask_events = (1ULL << (quant_bit * (uint32_t)e)) - 1, mask_l_quant = (1UL << lq) - 1;
	int full = 0;
	uint32_t buf_pos = 0, l_sigpos = 0, m = 0;
	uint64_t quant = 0;
	for (uint32_t i = 0; i < len; ++i) {
		if ((i > 0 && fabs(ev[i] - ev[l_sigpos]) < 0.3F) || ev[i] == 3.402823466e+32F) continue;   /* :243, LAST_SIG_DIFF, RI_MASK_SIGNAL */
		l_sigpos = i;
		uint32_t signal; memcpy(&signal, &ev[i], 4);                                                 /* :246 */
		uint32_t tq = (uint32_t)((signal >> 30 << lq) | ((signal >> shift_r) & mask_l_quant));       /* :247 */
		if (++buf_pos == (uint32_t)e) { full = 1; buf_pos = 0; }                                     /* :251 */
		quant = (quant << quant_bit | tq) & mask_events;                                             /* :254 */
		if (!full) continue;
		hash[m] = ora_hash64(quant, mask);                       /* x = hash << 6 | span (:255); the caller shifts it back (:372) */
		pos[m] = i;
		++m;
	}
	return m;
}

/* The index as a flat sorted table: keys[] unique hash values ascending, key_off[n_keys+1] into vals[]; vals of one key in
 * ri_idx_get's order (ascending y, rawindex.cpp:232).  y = ref_id << 32 | position << 1 | strand (rsketch.h:40). */
static inline int64_t ora_idx_find(const ora_index_t *ix, uint64_t h)
{
	int64_t lo = 0, hi = (int64_t)ix->n_keys - 1;
	while (lo <= hi) {
		int64_t mid = (lo + hi) >> 1;
		if (ix->keys[mid] < h) lo = mid + 1; else if (ix->keys[mid] > h) hi = mid - 1; else return mid;
	}
	return -1;
}
uint32_t ora_idx_get(const ora_index_t *ix, uint64_t h, const uint64_t **vals)
{
	int64_t k = ora_idx_find(ix, h);
	if (k < 0) { *vals = 0; return 0; }
	*vals = ix->vals + ix->key_off[k];
	return (uint32_t)(ix->key_off[k + 1] - ix->key_off[k]);
}

/* ================================================================================================
 * f1: gen_chains, rmap.cpp:315-543
 * ============================================================================================== */
typedef struct { ora_anchor_t *a; size_t n, m; } avec_t;
static void av_push(avec_t *v, ora_anchor_t x)
{
	if (v->n == v->m) { v->m = v->m ? v->m * 2 : 64; v->a = (ora_anchor_t*)realloc(v->a, v->m * sizeof(ora_anchor_t)); }
	v->a[v->n++] = x;
}
static int anchor_cmp(const void *pa, const void *pb)                    /* rmap.h:25-27: (target, query) ascending */
{
	const ora_anchor_t *a = (const ora_anchor_t*)pa, *b = (const ora_anchor_t*)pb;
	if (a->target_position != b->target_position) return a->target_position < b->target_position ? -1 : 1;
	if (a->query_position != b->query_position) return a->query_position < b->query_position ? -1 : 1;
	return 0;
}

typedef struct { ora_chain_t *c; size_t n, m; } cvec_t;
static void cv_push(cvec_t *v, ora_chain_t x)
{
	if (v->n == v->m) { v->m = v->m ? v->m * 2 : 16; v->c = (ora_chain_t*)realloc(v->c, v->m * sizeof(ora_chain_t)); }
	v->c[v->n++] = x;
}

struct ora_reader {
	const ora_index_t *ix;
	ora_mapopt_t opt;
	int contract;
	float *events; uint32_t n_events, cap_events;     /* p->events[read_id], rmap.cpp:554-567 */
	uint32_t offset;                                  /* reg->offset */
	ora_chain_t *chains; uint32_t n_chains;           /* reg->chains */
	cvec_t cand;                                      /* chains in evaluation order before the DTW filter (instrumentation) */
	uint32_t n_chunks_seen;
	uint64_t cells;                                   /* DP cells evaluated so far (instrumentation) */
	uint64_t n_seeds, n_hits;
};

static void free_chains(ora_chain_t *c, uint32_t n) { for (uint32_t i = 0; i < n; ++i) free(c[i].anchors); free(c); }

/* traceback_chains, rmap.cpp:130-173 */
static void traceback_chain(int min_num_anchors, int strand, size_t end_idx, uint32_t target_idx, const float *score, const size_t *pred,
                            const ora_anchor_t *anc, uint8_t *used, cvec_t *chains)
{
	if (used[end_idx]) return;
	avec_t v = {0, 0, 0};
	int stop_at_used = 0;
	size_t start = end_idx;
	av_push(&v, anc[start]);
	if (pred[start] != start && used[pred[start]]) stop_at_used = 1;
	used[start] = 1;
	uint32_t n = 1;
	while (pred[start] != start && !used[pred[start]]) {
		start = pred[start];
		av_push(&v, anc[start]);
		if (pred[start] != start && used[pred[start]]) stop_at_used = 1;
		used[start] = 1;
		++n;
	}
	if (n >= (uint32_t)min_num_anchors) {
		float adj = score[end_idx];
		if (stop_at_used) adj -= score[pred[start]];
		ora_chain_t c; memset(&c, 0, sizeof(c));
		c.chaining_score = adj; c.alignment_score = 0.0f; c.reference_sequence_index = target_idx;
		c.start_position = anc[start].target_position; c.end_position = anc[end_idx].target_position;
		c.n_anchors = n; c.mapq = 0; c.strand = strand; c.anchors = v.a;
		cv_push(chains, c);
	} else free(v.a);
}

typedef struct { float s; size_t i; } cand_t;
static int cand_cmp(const void *pa, const void *pb)                      /* compare(), rmap.cpp:175-179: score desc, index desc */
{
	const cand_t *a = (const cand_t*)pa, *b = (const cand_t*)pb;
	if (a->s > b->s) return -1;
	if (a->s < b->s) return 1;
	if (a->i > b->i) return -1;
	if (a->i < b->i) return 1;
	return 0;
}

/* std::sort with a strict "greater" comparator on <= 16 elements is libstdc++'s insertion sort, which never moves an
 * element past an equal one: a stable sort reproduces it.  Beyond 16 elements introsort's order of EQUAL keys is
 * unspecified (SURVEY 8c lists this as the one legitimate divergence); we stay stable there too. */
static void stable_sort_chains(ora_chain_t *c, size_t n, int (*greater)(const ora_chain_t*, const ora_chain_t*))
{
	for (size_t i = 1; i < n; ++i) {
		ora_chain_t x = c[i];
		size_t j = i;
		while (j > 0 && greater(&x, &c[j - 1])) { c[j] = c[j - 1]; --j; }
		c[j] = x;
	}
}
static int gt_chaining(const ora_chain_t *a, const ora_chain_t *b) { return a->chaining_score > b->chaining_score; }   /* rmap.cpp:512 */
static int gt_tuple(const ora_chain_t *a, const ora_chain_t *b)                                                         /* rmap.h:41-45 */
{
	if (a->alignment_score != b->alignment_score) return a->alignment_score > b->alignment_score;
	if (a->chaining_score != b->chaining_score) return a->chaining_score > b->chaining_score;
	if (a->n_anchors != b->n_anchors) return a->n_anchors > b->n_anchors;
	if (a->strand != b->strand) return a->strand > b->strand;
	if (a->reference_sequence_index != b->reference_sequence_index) return a->reference_sequence_index > b->reference_sequence_index;
	if (a->start_position != b->start_position) return a->start_position > b->start_position;
	return a->end_position > b->end_position;
}

/* gen_primary_chains, rmap.cpp:90-128 */
static void gen_primary(cvec_t *chains, const ora_mapopt_t *opt)
{
	stable_sort_chains(chains->c, chains->n, gt_tuple);
	const int dtw = (opt->flag & ORA_M_DTW_EVALUATE_CHAINS) != 0;
	cvec_t prim = {0, 0, 0};
	cv_push(&prim, chains->c[0]);
	size_t ci = 1;
	for (; ci < chains->n; ++ci) {
		const ora_chain_t *last = &prim.c[prim.n - 1];
		if (dtw) { if (chains->c[ci].alignment_score < last->alignment_score / 3) break; }
		else if (chains->c[ci].chaining_score < last->chaining_score / 3) break;
		int is_primary = 1;
		for (size_t pi = 0; pi < prim.n; ++pi)
			if (chains->c[ci].reference_sequence_index == prim.c[pi].reference_sequence_index) {
				uint32_t mx = chains->c[ci].start_position > prim.c[pi].start_position ? chains->c[ci].start_position : prim.c[pi].start_position;
				uint32_t mn = chains->c[ci].end_position < prim.c[pi].end_position ? chains->c[ci].end_position : prim.c[pi].end_position;
				if (mx <= mn) { is_primary = 0; break; }
			}
		if (is_primary) cv_push(&prim, chains->c[ci]); else free(chains->c[ci].anchors);
	}
	for (; ci < chains->n; ++ci) free(chains->c[ci].anchors);         /* the reference leaks these into the per-read pool (:100-104) */
	free(chains->c);
	*chains = prim;
}

/* comp_mapq, rmap.cpp:65-88 */
static void comp_mapq(cvec_t *chains, const ora_mapopt_t *opt)
{
	if (chains->n == 1) { chains->c[0].mapq = 60; return; }
	int mapq;
	if (opt->flag & ORA_M_DTW_EVALUATE_CHAINS) mapq = (int)(40 * (1 - chains->c[1].alignment_score / chains->c[0].alignment_score));
	else mapq = (int)(40 * (1 - chains->c[1].chaining_score / chains->c[0].chaining_score));
	if (mapq > 60) mapq = 60;
	if (mapq < 0) mapq = 0;
	chains->c[0].mapq = (uint8_t)mapq;
}

/* gen_chains for one chunk's events (chunk_events/l_chunk_events at query offset chunk_start). */
static void gen_chains(struct ora_reader *R, const float *chunk_events, uint32_t l_chunk_events, uint32_t chunk_start)
{
	const ora_index_t *ix = R->ix;
	const ora_mapopt_t *opt = &R->opt;
	const size_t n_seq = ix->n_seq;
	const int max_gap_length = (int)opt->max_gap_length, max_target_gap_length = (int)opt->max_target_gap_length;
	const int chaining_band_length = (int)opt->chaining_band_length, max_num_skips = (int)opt->max_num_skips;
	const int min_num_anchors = (int)opt->min_num_anchors, num_best_chains = (int)opt->num_best_chains;
	const float min_chaining_score = opt->min_chaining_score;
	const uint32_t mask = (uint32_t)((1ULL << 31) - 1);
	avec_t *fr[2];
	fr[0] = (avec_t*)calloc(n_seq, sizeof(avec_t)); fr[1] = (avec_t*)calloc(n_seq, sizeof(avec_t));

	for (uint32_t c = 0; c < R->n_chains; ++c)                        /* :343-357 anchors of the previous chunk's chains */
		for (uint32_t a = 0; a < R->chains[c].n_anchors; ++a) av_push(&fr[R->chains[c].strand][R->chains[c].reference_sequence_index], R->chains[c].anchors[a]);
	free_chains(R->chains, R->n_chains); R->chains = 0; R->n_chains = 0;   /* :359-362 */

	uint64_t *hash = (uint64_t*)malloc((size_t)(l_chunk_events + 1) * 8);
	uint32_t *pos = (uint32_t*)malloc((size_t)(l_chunk_events + 1) * 4);
	uint32_t n_seed = ora_sketch(chunk_events, l_chunk_events, ix->w, ix->e, ix->n, ix->q, ix->lq, ix->k, hash, pos, 0);   /* :367 */
	R->n_seeds += n_seed;
	for (uint32_t i = 0; i < n_seed; ++i) {                           /* :370-391 */
		const uint64_t *cr; uint32_t t = ora_idx_get(ix, hash[i], &cr);
		R->n_hits += t;
		for (uint32_t s = 0; s < t; ++s) {
			uint64_t kv = cr[s];
			uint32_t t_ind = (uint32_t)(kv >> 32), tpos = (uint32_t)(kv >> 1) & mask;
			ora_anchor_t a = { tpos, pos[i] + chunk_start };
			av_push(&fr[kv & 1][t_ind], a);
		}
	}
	free(hash); free(pos);
	for (size_t t = 0; t < n_seq; ++t) {                              /* :396-401 */
		if (fr[0][t].n) qsort(fr[0][t].a, fr[0][t].n, sizeof(ora_anchor_t), anchor_cmp);
		if (fr[1][t].n) qsort(fr[1][t].a, fr[1][t].n, sizeof(ora_anchor_t), anchor_cmp);
	}

	float max_chaining_score = 0;                                     /* :428, shared by every (target, strand) list */
	cvec_t chains = {0, 0, 0};
	for (size_t target = 0; target < n_seq; ++target)
		for (int strand = 0; strand < 2; ++strand) {
			const ora_anchor_t *anc = fr[strand][target].a;
			const size_t na = fr[strand][target].n;
			if (na == 0) continue;
			float *score = (float*)malloc(na * sizeof(float));
			size_t *pred = (size_t*)malloc(na * sizeof(size_t));
			uint8_t *used = (uint8_t*)calloc(na, 1);
			cand_t *cand = (cand_t*)malloc(na * sizeof(cand_t)); size_t n_cand = 0;
			for (size_t ai = 0; ai < na; ++ai) {                      /* :440-494 */
				score[ai] = (float)ix->e; pred[ai] = ai;
				int32_t ct = (int32_t)anc[ai].target_position, cq = (int32_t)anc[ai].query_position;
				int32_t start = 0;
				if (ai > (size_t)chaining_band_length) start = (int32_t)(ai - (size_t)chaining_band_length);
				int32_t num_skips = 0;
				for (int32_t pi = (int32_t)ai - 1; pi >= start; --pi) {
					int32_t pt = (int32_t)anc[pi].target_position, pq = (int32_t)anc[pi].query_position;
					if (pq == cq) continue;
					if (pt == ct) continue;
					if (pt + max_target_gap_length < ct) break;
					int32_t td = ct - pt, qd = cq - pq;
					float cur = 0;
					if (qd < 0) continue;
					int32_t md = td < qd ? td : qd; if (ix->e < md) md = ix->e;
					float matching = (float)md;
					int gap = abs(td - qd);
					float gap_scale = td > 0 ? (float)qd / (float)td : 1;
					if (gap < max_gap_length && gap_scale < 5 && gap_scale > 0.75) cur = score[pi] + matching;
					if (cur > score[ai]) { score[ai] = cur; pred[ai] = (size_t)pi; --num_skips; }
					else { ++num_skips; if (num_skips > max_num_skips) break; }
				}
				if (score[ai] > max_chaining_score) max_chaining_score = score[ai];
				if ((opt->flag & ORA_M_DISABLE_CHAININGSCORE_FILTERING) || (score[ai] >= min_chaining_score && score[ai] > max_chaining_score / 2)) {
					cand[n_cand].s = score[ai]; cand[n_cand].i = ai; ++n_cand;
				}
			}
			qsort(cand, n_cand, sizeof(cand_t), cand_cmp);           /* :496 */
			for (size_t k = 0; k < n_cand && k < (size_t)num_best_chains; ++k) {   /* :498-505 */
				traceback_chain(min_num_anchors, strand, cand[k].i, (uint32_t)target, score, pred, anc, used, &chains);
				if (!(opt->flag & ORA_M_DISABLE_CHAININGSCORE_FILTERING))
					if (score[cand[k].i] < max_chaining_score / 2) break;
			}
			free(score); free(pred); free(used); free(cand);
		}
	for (size_t t = 0; t < n_seq; ++t) { free(fr[0][t].a); free(fr[1][t].a); }
	free(fr[0]); free(fr[1]);

	/* candidates in evaluation order, for the workload generator / stage-wise GPU parity (copied, not moved) */
	for (size_t k = 0; k < R->cand.n; ++k) free(R->cand.c[k].anchors);
	R->cand.n = 0;

	if (opt->flag & (ORA_M_DTW_EVALUATE_CHAINS | ORA_M_DTW_LOG_SCORES)) {   /* :509-531 */
		stable_sort_chains(chains.c, chains.n, gt_chaining);          /* :512 */
		for (size_t k = 0; k < chains.n; ++k) {
			ora_chain_t c = chains.c[k];
			c.anchors = (ora_anchor_t*)malloc(c.n_anchors * sizeof(ora_anchor_t));
			memcpy(c.anchors, chains.c[k].anchors, c.n_anchors * sizeof(ora_anchor_t));
			cv_push(&R->cand, c);
		}
		ora_dtwopt_t d = { opt->dtw_border_constraint, opt->dtw_fill_method, opt->dtw_band_radius_frac, opt->dtw_match_bonus, opt->dtw_min_score };
		cvec_t post = {0, 0, 0};
		float best = 0.0f;                                            /* :515 */
		for (size_t k = 0; k < chains.n; ++k) {
			ora_chain_t *c = &chains.c[k];
			const float *ref = c->strand == 1 ? ix->fwd[c->reference_sequence_index] : ix->rev[c->reference_sequence_index];   /* :183-188 */
			c->alignment_score = ora_align_chain(c->anchors, c->n_anchors, R->events, ref, &d, best, 1, &R->cells);
			R->cand.c[k].alignment_score = c->alignment_score;
			if (c->alignment_score >= opt->dtw_min_score) {
				if (c->alignment_score > best) best = c->alignment_score;
				cv_push(&post, *c);
			} else if (opt->flag & ORA_M_DTW_EVALUATE_CHAINS) free(c->anchors);
		}
		if (opt->flag & ORA_M_DTW_EVALUATE_CHAINS) { free(chains.c); chains = post; }
		else free(post.c);
	}
	if (chains.n > 0) {                                               /* :533-542 */
		gen_primary(&chains, opt);
		comp_mapq(&chains, opt);
		R->n_chains = (uint32_t)chains.n; R->chains = chains.c;
	} else free(chains.c);
}

/* is_mapped_with_high_confidence, rmap.cpp:594-665 */
static int is_mapped(const struct ora_reader *R)
{
	const ora_mapopt_t *opt = &R->opt;
	uint32_t n_anchors0 = R->n_chains ? R->chains[0].n_anchors : 0;
	if (n_anchors0 == 0) return 0;
	const int dtw = (opt->flag & ORA_M_DTW_EVALUATE_CHAINS) != 0;
#define SC(c) (dtw ? (c).alignment_score : (c).chaining_score)
	if (R->n_chains >= 2) {
		if (SC(R->chains[0]) / SC(R->chains[1]) >= opt->min_bestmap_ratio) return 1;
		float mean = 0;
		for (uint32_t c = 0; c < R->n_chains; ++c) mean += SC(R->chains[c]);
		mean /= R->n_chains;
		if (SC(R->chains[0]) >= opt->min_meanmap_ratio * mean) return 1;
	} else if (R->n_chains == 1 && R->chains[0].n_anchors >= opt->min_chain_anchor) return 1;
#undef SC
	return 0;
}

struct ora_reader *ora_reader_new(const ora_index_t *ix, const ora_mapopt_t *opt, int contract)
{
	struct ora_reader *R = (struct ora_reader*)calloc(1, sizeof(*R));
	R->ix = ix; R->opt = *opt; R->contract = contract;
	return R;
}
void ora_reader_free(struct ora_reader *R)
{
	if (!R) return;
	free(R->events); free_chains(R->chains, R->n_chains);
	for (size_t k = 0; k < R->cand.n; ++k) free(R->cand.c[k].anchors);
	free(R->cand.c); free(R);
}

/* ri_map_frag, rmap.cpp:545-578, on events already detected (so the GPU's events can be fed to the same code) */
int ora_reader_chunk_events(struct ora_reader *R, const float *cev, uint32_t n_ev)
{
	if (R->n_events + n_ev > R->cap_events) { R->cap_events = (R->n_events + n_ev) * 2 + 64; R->events = (float*)realloc(R->events, (size_t)R->cap_events * 4); }
	memcpy(R->events + R->n_events, cev, (size_t)n_ev * 4); R->n_events += n_ev;   /* :554-567 */
	R->n_chunks_seen++;
	for (size_t k = 0; k < R->cand.n; ++k) free(R->cand.c[k].anchors);
	R->cand.n = 0;
	if (n_ev < R->opt.min_events) return is_mapped(R);                /* :569-572: no chaining, offset not advanced */
	gen_chains(R, cev, n_ev, R->offset);                              /* :574 */
	R->offset += n_ev;                                                /* :575 */
	return is_mapped(R);
}
int ora_reader_chunk(struct ora_reader *R, const float *sig, uint32_t s_len)
{
	ora_evopt_t eo = { R->opt.window_length1, R->opt.window_length2, R->opt.threshold1, R->opt.threshold2, R->opt.peak_height };
	float *ev = (float*)malloc((size_t)(s_len + 1) * 4);
	uint32_t *pk = (uint32_t*)malloc((size_t)(s_len + 1) * 4);
	uint32_t npk = 0;
	uint32_t n_ev = ora_detect_events(sig, s_len, &eo, R->contract, ev, pk, &npk);   /* :551 */
	int m = ora_reader_chunk_events(R, ev, n_ev);
	free(ev); free(pk);
	return m;
}
uint32_t ora_reader_state(const struct ora_reader *R, uint32_t *n_events, uint32_t *offset, uint64_t *cells, uint64_t *n_seeds, uint64_t *n_hits)
{
	if (n_events) *n_events = R->n_events;
	if (offset) *offset = R->offset;
	if (cells) *cells = R->cells;
	if (n_seeds) *n_seeds = R->n_seeds;
	if (n_hits) *n_hits = R->n_hits;
	return R->n_chains;
}
const float *ora_reader_events(const struct ora_reader *R) { return R->events; }
/* which = 0: reg->chains after the chunk; which = 1: the candidates in evaluation order before the DTW filter */
uint32_t ora_reader_chains(const struct ora_reader *R, int which, ora_chain_out_t *out, uint32_t cap, ora_anchor_t *anchors_out, uint64_t acap)
{
	const ora_chain_t *c = which ? R->cand.c : R->chains;
	uint32_t n = which ? (uint32_t)R->cand.n : R->n_chains;
	uint64_t ao = 0;
	for (uint32_t k = 0; k < n && k < cap; ++k) {
		out[k].chaining_score = c[k].chaining_score; out[k].alignment_score = c[k].alignment_score;
		out[k].reference_sequence_index = c[k].reference_sequence_index; out[k].start_position = c[k].start_position;
		out[k].end_position = c[k].end_position; out[k].n_anchors = c[k].n_anchors; out[k].mapq = c[k].mapq; out[k].strand = c[k].strand;
		out[k].anchor_off = ao;
		for (uint32_t a = 0; a < c[k].n_anchors && ao < acap; ++a) anchors_out[ao++] = c[k].anchors[a];
	}
	return n;
}

/* ================================================================================================
 * map_worker_for, rmap.cpp:667-822: the chunk loop and the PAF fields of one read
 * ============================================================================================== */
void ora_map_read(const ora_index_t *ix, const ora_mapopt_t *opt, int contract, const float *sig, uint32_t qlen, ora_paf_t *out)
{
	struct ora_reader *R = ora_reader_new(ix, opt, contract);
	const uint32_t l_chunk = opt->chunk_size, max_chunk = opt->max_num_chunk;
	uint32_t chunk_start, current_chunk;
	for (chunk_start = current_chunk = 0; chunk_start < qlen && current_chunk < max_chunk; chunk_start += l_chunk, ++current_chunk) {   /* :685 */
		uint32_t chunk_end = chunk_start + l_chunk;
		if (chunk_end > qlen) chunk_end = qlen;
		if (ora_reader_chunk(R, sig + chunk_start, chunk_end - chunk_start)) break;                                                     /* :689-692 */
	}
	ora_paf_finish(R, qlen, chunk_start, current_chunk, out);
	ora_reader_free(R);
}

/* rmap.cpp:696-800 given the loop's exit state.  aln (cigar) is produced separately by the caller. */
void ora_paf_finish(const struct ora_reader *R, uint32_t qlen, uint32_t chunk_start, uint32_t current_chunk, ora_paf_t *o)
{
	const ora_mapopt_t *opt = &R->opt;
	const uint32_t l_chunk = opt->chunk_size, max_chunk = opt->max_num_chunk;
	memset(o, 0, sizeof(*o));
	if (current_chunk > 0 && (chunk_start >= qlen || current_chunk == max_chunk)) --current_chunk;                                      /* :696 */
	float read_position_scale = ((float)(current_chunk + 1) * l_chunk / R->offset) / ((float)opt->sample_rate / opt->bp_per_sec);        /* :698 */
	const ora_chain_t *chains = R->chains;
	uint32_t n_chains = chains ? R->n_chains : 0;
	uint32_t n_anchors0 = n_chains ? chains[0].n_anchors : 0;
	float mean_chain_score = 0;
	for (uint32_t c = 0; c < n_chains; ++c) mean_chain_score += chains[c].chaining_score;
	mean_chain_score /= n_chains;                                                                                                       /* :709 (0/0 = nan when no chains; unused then) */
	o->ci = current_chunk + 1; o->sl = qlen; o->n_events = R->n_events; o->cells = R->cells;
	float at = 0, aq = 0;
	if (n_chains >= 1) {
		for (size_t ai = 0; ai + 1 < n_anchors0; ++ai) {
			at += chains[0].anchors[ai].target_position - chains[0].anchors[ai + 1].target_position;
			aq += chains[0].anchors[ai].query_position - chains[0].anchors[ai + 1].query_position;
		}
	}
	if (is_mapped(R)) {
		at /= n_anchors0; aq /= n_anchors0;
		o->mapped = 1; o->cm = n_anchors0; o->nc = n_chains; o->s1 = chains[0].chaining_score; o->s2 = n_chains > 1 ? chains[1].chaining_score : 0;
		o->sm = mean_chain_score; o->at = at; o->aq = aq; o->alns = chains[0].alignment_score;
		o->ref_id = chains[0].reference_sequence_index;
		o->read_length = (uint32_t)(read_position_scale * chains[0].anchors[0].query_position);
		o->read_start_position = (uint32_t)(read_position_scale * chains[0].anchors[n_anchors0 - 1].query_position);
		o->read_end_position = (uint32_t)(read_position_scale * chains[0].anchors[0].query_position);
		o->fragment_start_position = chains[0].strand ? (uint32_t)(R->ix->seq_len[chains[0].reference_sequence_index] + 1 - chains[0].end_position) : chains[0].start_position;
		o->fragment_length = (uint32_t)(chains[0].end_position - chains[0].start_position + 1);
		o->mapq = chains[0].mapq & 63; o->rev = chains[0].strand & 1;                                                                  /* mapq is a 6-bit field, rmap.h:57 */
	} else {
		if (n_chains >= 1) {
			if (n_anchors0) { at /= n_anchors0; aq /= n_anchors0; }
			o->cm = n_anchors0; o->nc = n_chains; o->s1 = chains[0].chaining_score; o->s2 = n_chains > 1 ? chains[1].chaining_score : 0;
			o->sm = mean_chain_score; o->at = at; o->aq = aq;
		}
		o->read_length = (uint32_t)(read_position_scale * R->offset);
	}
}

/* The PAF line of rmap.cpp:954-972 without the mt:f: tag (per-read wall time).  Returns the length. */
int ora_paf_format(const ora_paf_t *o, const char *read_name, const char *ref_name, uint32_t ref_len, char *buf, int cap)
{
	int n;
	if (o->mapped)
		n = snprintf(buf, cap, "%s\t%u\t%u\t%u\t%c\t%s\t%u\t%u\t%u\t%u\t%u\t%d\tci:i:%u\tsl:i:%u\tcm:i:%u\tnc:i:%u\ts1:f:%f\ts2:f:%f\tsm:f:%f\tat:f:%f\taq:f:%f",
		             read_name, o->read_length, o->read_start_position, o->read_end_position, o->rev ? '-' : '+', ref_name, ref_len,
		             o->fragment_start_position, o->fragment_start_position + o->fragment_length, o->read_end_position - o->read_start_position - 1,
		             o->fragment_length, (int)o->mapq, o->ci, o->sl, o->cm, o->nc, o->s1, o->s2, o->sm, o->at, o->aq);
	else if (o->nc >= 1)
		n = snprintf(buf, cap, "%s\t%u\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\tci:i:%u\tsl:i:%u\tcm:i:%u\tnc:i:%u\ts1:f:%f\ts2:f:%f\tsm:f:%f\tat:f:%f\taq:f:%f",
		             read_name, o->read_length, (int)o->mapq, o->ci, o->sl, o->cm, o->nc, o->s1, o->s2, o->sm, o->at, o->aq);
	else
		n = snprintf(buf, cap, "%s\t%u\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\tci:i:%u\tsl:i:%u\tcm:i:0\tnc:i:0\ts1:f:0\ts2:f:0\tsm:f:0\tat:f:0\taq:f:0",
		             read_name, o->read_length, (int)o->mapq, o->ci, o->sl);
	return n;
}
