/*
 * refmap_shim.cpp -- C entry points around the UNMODIFIED reference map loop.  TEST INFRASTRUCTURE.
 *
 * oracle/Makefile (target `refmap`) compiles this translation unit with the reference's own C++ flags.
 * It #includes /root/reference/src/rmap.cpp where it lies (nothing is copied into this repository), so
 * the file-local functions of the map loop -- align_chain (rmap.cpp:181), gen_chains (:315),
 * ri_map_frag (:545), is_mapped_with_high_confidence (:594), map_worker_for (:667) -- are reachable,
 * and links the reference's other objects (dtw, revent, rsketch, rawindex, rsig, pore_model, ...) and its
 * vendored HDF5 into oracle/_ref/libra_refmap.so.  The shim adds no arithmetic of its own: it unpacks
 * arguments, calls the reference function and copies the result out.  The only loops restated here are
 * the 8-line evaluation loop of rmap.cpp:515-524 (it sits in the middle of gen_chains) around the real
 * align_chain, and the PAF fprintf of rmap.cpp:954-972 (it prints to stdout) as an snprintf.
 *
 * Also here (our code, not the reference's): a small multi-read FAST5 writer on the HDF5 C API, so that
 * the reference's reader (rsig.cpp:43-96,186-229) and the full `rawalign` binary can be driven with
 * synthetic reads.
 */
#include "rmap.cpp"          /* -I/root/reference/src */
#include "pore_model.h"
#include <hdf5.h>
#include <omp.h>
#include <string>
#include <vector>

extern "C" {

/* ---------------------------------------------------------------- options ------------------------ */
uint32_t refmap_sizeof_mapopt(void) { return (uint32_t)sizeof(ri_mapopt_t); }
void refmap_mapopt_init(ri_mapopt_t *o) { ri_mapopt_init(o); }

/* ---------------------------------------------------------------- index -------------------------- */
int refmap_pore_model(const char *path, float *level_mean, uint32_t cap)
{
	sigmap::PoreModel pm;
	pm.Load(std::string(path));
	int k = pm.GetKmerSize();
	uint32_t n = 1U << (2 * k);
	for (uint32_t i = 0; i < n && i < cap; ++i) level_mean[i] = pm.pore_models_[i].level_mean;   /* main.cpp:345-347 */
	return k;
}

/* main.cpp:323-354: open, (load pore model), ri_idx_reader_read; optional dump (-d). */
void *refmap_index_build(const char *fasta, const char *model, int e, int q, int lq, int w, int n, int k, int n_threads,
                         const char *dump_path)
{
	ri_idxopt_t ipt; ri_idxopt_init(&ipt);
	ipt.e = e; ipt.q = q; ipt.lq = lq; ipt.w = w; ipt.n = n; ipt.k = k;
	ri_idx_reader_t *r = ri_idx_reader_open(fasta, &ipt, dump_path);
	if (!r) return 0;
	float *pore_vals = 0;
	if (!r->is_idx) {
		sigmap::PoreModel pm; pm.Load(std::string(model));
		pore_vals = (float*)calloc(1U << (pm.GetKmerSize() * 2), sizeof(float));
		for (uint32_t i = 0; i < 1U << (pm.GetKmerSize() * 2); ++i) pore_vals[i] = pm.pore_models_[i].level_mean;
	}
	ri_idx_t *ri = ri_idx_reader_read(r, pore_vals, n_threads);
	ri_idx_reader_close(r);
	if (pore_vals) free(pore_vals);
	return ri;
}
void *refmap_index_load(const char *path) { return refmap_index_build(path, 0, 0, 0, 0, 0, 0, 0, 1, 0); }
void refmap_index_free(void *ri) { if (ri) ri_idx_destroy((ri_idx_t*)ri); }
void refmap_index_info(void *h, int32_t *out)
{
	ri_idx_t *ri = (ri_idx_t*)h;
	out[0] = ri->b; out[1] = ri->w; out[2] = ri->e; out[3] = ri->n; out[4] = ri->q; out[5] = ri->lq; out[6] = ri->k;
	out[7] = ri->flag; out[8] = (int32_t)ri->n_seq;
}
uint32_t refmap_index_seq(void *h, uint32_t i, char *name, uint32_t cap)
{
	ri_idx_t *ri = (ri_idx_t*)h;
	if (name && cap) { strncpy(name, ri->seq[i].name ? ri->seq[i].name : "", cap - 1); name[cap - 1] = 0; }
	return ri->seq[i].len;
}
const float *refmap_index_signal(void *h, uint32_t i, int forward)
{
	ri_idx_t *ri = (ri_idx_t*)h;
	return forward ? ri->forward_signals[i] : ri->reverse_signals[i];
}
int refmap_index_get(void *h, uint64_t hashval, uint64_t *out, int cap)
{
	int t = 0;
	const uint64_t *cr = ri_idx_get((ri_idx_t*)h, hashval, &t);
	for (int s = 0; s < t && s < cap; ++s) out[s] = cr[s];
	return t;
}

/* ---------------------------------------------------------------- stages ------------------------- */
uint32_t refmap_sketch(const float *ev, int len, int w, int e, int n, int q, int lq, int k, uint64_t *x, uint64_t *y, uint32_t cap)
{
	mm128_v riv = {0, 0, 0};
	ri_sketch(0, ev, 0, 0, len, w, e, n, q, lq, k, &riv);        /* exactly the call of rmap.cpp:367 */
	uint32_t m = (uint32_t)riv.n;
	for (uint32_t i = 0; i < m && i < cap; ++i) { x[i] = riv.a[i].x; y[i] = riv.a[i].y; }
	ri_kfree(0, riv.a);
	return m;
}
void refmap_seq_to_sig(const char *seq, int len, const float *pore_vals, int k, int strand, uint32_t *s_len, float *out)
{ ri_seq_to_sig(seq, len, pore_vals, k, strand, s_len, out); }

/* ---------------------------------------------------------------- align_chain -------------------- */
static inline void one_ref_idx(ri_idx_t *ri, float **slot, const float *ref)
{
	memset(ri, 0, sizeof(*ri));
	slot[0] = (float*)ref;
	ri->forward_signals = slot; ri->reverse_signals = slot; ri->n_seq = 1;
}

/* One call of align_chain (rmap.cpp:181-313).  ref = the strand's whole reference signal, read = the read's whole
 * event array.  With cigar != 0 the alignment (rmap.cpp:226-235, :285-299) is copied out. */
float refmap_align_chain(const ri_anchor_t *anchors, uint32_t n_anchors, const float *read, const float *ref,
                         const ri_mapopt_t *opt, int cigar, float min_score, float *dtw_cost,
                         uint64_t *pi, uint64_t *pj, float *pd, uint64_t cap, uint64_t *n_path)
{
	ri_idx_t ri; float *slot[1]; one_ref_idx(&ri, slot, ref);
	ri_chain_t ch; memset((void*)&ch, 0, sizeof(ch));
	ch.n_anchors = n_anchors; ch.anchors = (ri_anchor_t*)anchors; ch.strand = 1; ch.reference_sequence_index = 0;
	if (!cigar) {
		align_chain(ch, &ri, read, 0, 0, opt, false, min_score);
		return ch.alignment_score;
	}
	align_chain(ch, &ri, read, 0, 0, opt, true);
	uint64_t m = ch.dtw_result.alignment.size();
	if (n_path) *n_path = m;
	if (dtw_cost) *dtw_cost = ch.dtw_result.cost;
	for (uint64_t k = 0; k < m && k < cap; ++k) {
		pi[k] = ch.dtw_result.alignment[k].position.i; pj[k] = ch.dtw_result.alignment[k].position.j;
		if (pd) pd[k] = ch.dtw_result.alignment[k].difference;
	}
	return ch.alignment_score;   /* ch's own destructor releases dtw_result */
}

/* same layout as ra_chain_t (include/rawalign_b200.h) */
typedef struct { uint64_t anchor_off; uint32_t n_anchors; uint32_t group; uint64_t read_off; uint64_t ref_off; } shim_chain_t;

/* rmap.cpp:515-524 around the real align_chain for the chains [lo, hi) of one group. */
static inline void eval_group(const float *read_arena, const float *ref_arena, const ri_anchor_t *anchors, const shim_chain_t *chains,
                              uint32_t lo, uint32_t hi, const ri_mapopt_t *opt, int running, const float *min_score,
                              float *score, uint8_t *keep)
{
	float best_found_alignment = 0.0f;
	for (uint32_t k = lo; k < hi; ++k) {
		const shim_chain_t &c = chains[k];
		ri_idx_t ri; float *slot[1]; one_ref_idx(&ri, slot, ref_arena + c.ref_off);
		ri_chain_t ch; memset((void*)&ch, 0, sizeof(ch));
		ch.n_anchors = c.n_anchors; ch.anchors = (ri_anchor_t*)(anchors + c.anchor_off); ch.strand = 1;
		float ms = running ? best_found_alignment : (min_score ? min_score[k] : -1e10f);
		align_chain(ch, &ri, read_arena + c.read_off, 0, 0, opt, false, ms);
		score[k] = ch.alignment_score;
		bool kp = ch.alignment_score >= opt->dtw_min_score;
		if (kp && ch.alignment_score > best_found_alignment) best_found_alignment = ch.alignment_score;
		if (keep) keep[k] = kp;
	}
}

/* Batched chain evaluation, all loops in C; groups are dealt to n_threads OpenMP threads dynamically (kt_for deals reads
 * the same way, kthread.c:24-52).  group_first has n_groups+1 entries. */
void refmap_evaluate_chains_batch(const float *read_arena, const float *ref_arena, const ri_anchor_t *anchors,
                                  const shim_chain_t *chains, const uint32_t *group_first, uint32_t n_groups,
                                  const ri_mapopt_t *opt, int running, const float *min_score, float *score, uint8_t *keep,
                                  int n_threads)
{
	#pragma omp parallel for schedule(dynamic, 4) num_threads(n_threads > 0 ? n_threads : 1)
	for (long g = 0; g < (long)n_groups; ++g)
		eval_group(read_arena, ref_arena, anchors, chains, group_first[g], group_first[g + 1], opt, running, min_score, score, keep);
}

/* The hot path of one batch of reads exactly as the reference's map loop runs it, minus seeding and chaining (those are
 * given): per read, per chunk: detect_events (rmap.cpp:551) and the append of rmap.cpp:554-567; then the chain groups of
 * that chunk round (rmap.cpp:515-524).  Chains address the read's events relative to the read (read_off is ignored here;
 * the per-read event array is the one this function builds).  round_first[r*(max_rounds+1) + c] .. [.. + c + 1] are the
 * groups' chain ranges.  ev_out/ev_count (optional): the read's events are copied to ev_out + ev_base[r].
 * Returns the number of events. */
uint64_t refmap_hotpath_batch(const float *sig, const uint64_t *read_sig_off, const uint32_t *read_l_sig, uint32_t n_reads,
                              uint32_t chunk_size, uint32_t max_chunks, const float *ref_arena, const ri_anchor_t *anchors,
                              const shim_chain_t *chains, const uint32_t *round_first /* n_reads*(max_chunks+1) */,
                              const ri_mapopt_t *opt, float *score, uint8_t *keep, float *ev_out, const uint64_t *ev_base,
                              uint32_t *ev_count, int n_threads)
{
	uint64_t total = 0;
	#pragma omp parallel num_threads(n_threads > 0 ? n_threads : 1) reduction(+:total)
	{
		void *km = ri_km_init();
		#pragma omp for schedule(dynamic, 1)
		for (long r = 0; r < (long)n_reads; ++r) {
			ri_events_t ge; ge.values = 0; ge.length = 0;
			const float *s = sig + read_sig_off[r];
			uint32_t qlen = read_l_sig[r], cur = 0;
			for (uint32_t cs = 0; cs < qlen && cur < max_chunks; cs += chunk_size, ++cur) {
				uint32_t ce = cs + chunk_size; if (ce > qlen) ce = qlen;
				uint32_t n_ev = 0;
				float *cev = detect_events(km, ce - cs, s + cs, opt, &n_ev);
				if (!ge.values) {                                                 /* rmap.cpp:554-567 */
					ge.values = (float*)malloc(n_ev * sizeof(float)); ge.length = n_ev;
					std::copy(cev, cev + n_ev, ge.values);
				} else {
					float *nw = (float*)malloc((ge.length + n_ev) * sizeof(float));
					std::copy(ge.values, ge.values + ge.length, nw);
					std::copy(cev, cev + n_ev, nw + ge.length);
					free(ge.values); ge.values = nw; ge.length += n_ev;
				}
				if (cev) ri_kfree(km, cev);
				if (chains) {
					uint32_t lo = round_first[r * (uint64_t)(max_chunks + 1) + cur], hi = round_first[r * (uint64_t)(max_chunks + 1) + cur + 1];
					if (hi > lo) {
						/* the chains carry read-relative offsets: read_arena = this read's array, read_off forced to 0 */
						float best_found_alignment = 0.0f;
						for (uint32_t k = lo; k < hi; ++k) {
							const shim_chain_t &c = chains[k];
							ri_idx_t ri; float *slot[1]; one_ref_idx(&ri, slot, ref_arena + c.ref_off);
							ri_chain_t ch; memset((void*)&ch, 0, sizeof(ch));
							ch.n_anchors = c.n_anchors; ch.anchors = (ri_anchor_t*)(anchors + c.anchor_off); ch.strand = 1;
							align_chain(ch, &ri, ge.values, ge.length, 0, opt, false, best_found_alignment);
							score[k] = ch.alignment_score;
							bool kp = ch.alignment_score >= opt->dtw_min_score;
							if (kp && ch.alignment_score > best_found_alignment) best_found_alignment = ch.alignment_score;
							if (keep) keep[k] = kp;
						}
					}
				}
			}
			if (ev_out && ge.values) std::copy(ge.values, ge.values + ge.length, ev_out + ev_base[r]);
			if (ev_count) ev_count[r] = ge.length;
			total += ge.length;
			free(ge.values);
		}
		ri_km_destroy(km);
	}
	return total;
}

/* ---------------------------------------------------------------- gen_chains, one read ------------- */
typedef struct {
	pipeline_mt *p; ri_reg1_t *reg; ri_tbuf_t *b; ri_mapopt_t opt; const ri_idx_t *ri; uint32_t n_chunks;
} reader_t;

void *refmap_reader_new(void *idx, const ri_mapopt_t *opt)
{
	reader_t *h = new reader_t();
	h->p = new pipeline_mt(); h->p->events.resize(1);
	h->p->n_processed = 0; h->p->n_threads = 1; h->p->opt = 0; h->p->ri = (const ri_idx_t*)idx;
	h->p->events[0].values = 0; h->p->events[0].length = 0; h->p->events[0].rid = 0; h->p->events[0].name = 0;
	h->reg = (ri_reg1_t*)calloc(1, sizeof(ri_reg1_t)); h->reg->read_id = 0; h->reg->read_name = "r";
	h->b = ri_tbuf_init(); h->opt = *opt; h->p->opt = &h->opt; h->ri = (const ri_idx_t*)idx; h->n_chunks = 0;
	return h;
}
/* ri_map_frag (rmap.cpp:545-578) for the next chunk; returns is_mapped_with_high_confidence (rmap.cpp:692). */
int refmap_reader_chunk(void *hh, const float *sig, uint32_t s_len)
{
	reader_t *h = (reader_t*)hh;
	ri_map_frag(h->p, h->ri, s_len, sig, h->reg, h->b, &h->opt, "r");
	h->n_chunks++;
	return is_mapped_with_high_confidence(h->reg, &h->opt) ? 1 : 0;
}
typedef struct {
	float chaining_score, alignment_score; uint32_t reference_sequence_index, start_position, end_position, n_anchors;
	uint32_t mapq; int32_t strand; uint64_t anchor_off;
} shim_chain_out_t;
/* reg->chains after the last chunk; anchors are concatenated into anchors_out.  Returns n_chains. */
uint32_t refmap_reader_chains(void *hh, shim_chain_out_t *out, uint32_t cap, ri_anchor_t *anchors_out, uint64_t acap,
                              uint32_t *n_events, uint32_t *reg_offset)
{
	reader_t *h = (reader_t*)hh;
	uint64_t ao = 0;
	for (uint32_t c = 0; c < h->reg->n_chains && c < cap; ++c) {
		const ri_chain_t &ch = h->reg->chains[c];
		out[c].chaining_score = ch.chaining_score; out[c].alignment_score = ch.alignment_score;
		out[c].reference_sequence_index = ch.reference_sequence_index; out[c].start_position = ch.start_position;
		out[c].end_position = ch.end_position; out[c].n_anchors = ch.n_anchors; out[c].mapq = ch.mapq; out[c].strand = ch.strand;
		out[c].anchor_off = ao;
		for (uint32_t a = 0; a < ch.n_anchors && ao < acap; ++a) anchors_out[ao++] = ch.anchors[a];
	}
	if (n_events) *n_events = h->p->events[0].length;
	if (reg_offset) *reg_offset = h->reg->offset;
	return h->reg->n_chains;
}
uint32_t refmap_reader_events(void *hh, float *out, uint32_t cap)
{
	reader_t *h = (reader_t*)hh;
	uint32_t n = h->p->events[0].length;
	for (uint32_t i = 0; i < n && i < cap; ++i) out[i] = h->p->events[0].values[i];
	return n;
}
void refmap_reader_free(void *hh)
{
	reader_t *h = (reader_t*)hh;
	free(h->p->events[0].values);
	ri_tbuf_destroy(h->b); free(h->reg); delete h->p; delete h;
}

/* ---------------------------------------------------------------- whole reads: map_worker_for ------ */
/* kt_for(n_threads, map_worker_for, ...) over a batch of already decoded reads (rmap.cpp:916), then the PAF lines of
 * rmap.cpp:954-972 written with snprintf instead of fprintf(stdout).  Lines are separated by '\n'.  Returns the number of
 * bytes written (truncated at cap). */
uint64_t refmap_map_reads(void *idx, const ri_mapopt_t *opt, const float *sig, const uint64_t *sig_off, const uint32_t *l_sig,
                          uint32_t n_reads, int n_threads, char *paf, uint64_t cap, uint8_t *mapped_out, uint32_t *chunks_out)
{
	const ri_idx_t *ri = (const ri_idx_t*)idx;
	pipeline_mt pl;
	pl.n_processed = 0; pl.n_threads = n_threads > 1 ? n_threads : 1; pl.opt = opt; pl.ri = ri; pl.su_stop = 0;
	pl.events.reserve(n_reads);
	step_mt s; memset((void*)&s, 0, sizeof(s));
	s.p = &pl; s.n_sig = (int)n_reads;
	s.sig = (ri_sig_t**)calloc(n_reads, sizeof(ri_sig_t*));
	s.reg = (ri_reg1_t**)calloc(n_reads, sizeof(ri_reg1_t*));
	std::vector<std::string> names(n_reads);
	for (uint32_t i = 0; i < n_reads; ++i) {
		names[i] = "read_" + std::to_string(i);
		s.sig[i] = (ri_sig_t*)calloc(1, sizeof(ri_sig_t));
		s.sig[i]->rid = pl.n_processed++; s.sig[i]->l_sig = l_sig[i]; s.sig[i]->name = (char*)names[i].c_str();
		s.sig[i]->sig = (float*)(sig + sig_off[i]);
		ri_events_t ev; ev.rid = s.sig[i]->rid; ev.name = s.sig[i]->name; ev.values = NULL; ev.length = 0;
		pl.events.push_back(ev);
		s.reg[i] = (ri_reg1_t*)calloc(1, sizeof(ri_reg1_t));
		s.reg[i]->read_id = s.sig[i]->rid; s.reg[i]->read_name = s.sig[i]->name;
	}
	s.buf = (ri_tbuf_t**)calloc(pl.n_threads, sizeof(ri_tbuf_t*));
	for (int i = 0; i < pl.n_threads; ++i) s.buf[i] = ri_tbuf_init();
	kt_for(pl.n_threads, map_worker_for, &s, s.n_sig);
	uint64_t w = 0;
	for (uint32_t k = 0; k < n_reads; ++k) {
		ri_reg1_t *reg0 = s.reg[k];
		char strand = reg0->rev ? '-' : '+';
		char line[1 << 16]; int n = 0;
		if (reg0->ref_id < ri->n_seq && reg0->read_name) {
			if (reg0->mapped)
				n = snprintf(line, sizeof(line), "%s\t%u\t%u\t%u\t%c\t%s\t%u\t%u\t%u\t%u\t%u\t%d\t%s\n",
				             reg0->read_name, reg0->read_length, reg0->read_start_position, reg0->read_end_position,
				             strand, ri->seq[reg0->ref_id].name, ri->seq[reg0->ref_id].len, reg0->fragment_start_position,
				             reg0->fragment_start_position + reg0->fragment_length, reg0->read_end_position - reg0->read_start_position - 1,
				             reg0->fragment_length, reg0->mapq, reg0->tags);
			else
				n = snprintf(line, sizeof(line), "%s\t%u\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\t%s\n", reg0->read_name, reg0->read_length, reg0->mapq, reg0->tags);
		}
		if (n > (int)sizeof(line) - 1) n = sizeof(line) - 1;
		if (paf && w + n < cap) { memcpy(paf + w, line, n); w += n; }
		if (mapped_out) mapped_out[k] = reg0->mapped;
		if (chunks_out) { const char *ci = reg0->tags ? strstr(reg0->tags, "ci:i:") : 0; chunks_out[k] = ci ? (uint32_t)atoi(ci + 5) : 0; }
		if (reg0->tags) free(reg0->tags);
		free(reg0);
	}
	if (paf && w < cap) paf[w] = 0;
	for (int i = 0; i < pl.n_threads; ++i) ri_tbuf_destroy(s.buf[i]);
	free(s.buf); free(s.reg);
	for (uint32_t i = 0; i < n_reads; ++i) { free(pl.events[i].values); free(s.sig[i]); }
	free(s.sig);
	return w;
}

/* ---------------------------------------------------------------- FAST5 ----------------------------- */
/* Reads every read of one FAST5 through the reference's reader (open_sig + ri_read_sig, rsig.cpp:113,186-229).
 * Pass sig_out == NULL to size: returns the number of reads and fills l_sig (cap_reads entries). */
int refmap_read_fast5(const char *path, float *sig_out, uint64_t sig_cap, uint32_t *l_sig, uint32_t cap_reads,
                      float *dig, float *ran, float *offset, char *names, uint32_t name_stride)
{
	ri_sig_file_t *fp = open_sig(path);
	if (!fp) return -1;
	int n = 0; uint64_t o = 0;
	while (fp->cur_read < fp->num_read) {
		ri_sig_t s; memset(&s, 0, sizeof(s));
		ri_read_sig(fp, &s);
		if ((uint32_t)n < cap_reads) {
			l_sig[n] = s.l_sig;
			if (dig) dig[n] = s.dig; if (ran) ran[n] = s.ran; if (offset) offset[n] = s.offset;
			if (names) { strncpy(names + (size_t)n * name_stride, s.name, name_stride - 1); names[(size_t)n * name_stride + name_stride - 1] = 0; }
			if (sig_out && o + s.l_sig <= sig_cap) memcpy(sig_out + o, s.sig, (size_t)s.l_sig * sizeof(float));
			o += s.l_sig;
		}
		free(s.sig); free(s.name);
		++n;
	}
	ri_sig_close(fp);
	return n;
}

/* Our writer (HDF5 C API): multi-read FAST5 with exactly the fields the reader touches:
 *   /<name>/Raw            attribute read_id (string), dataset Signal (int16)
 *   /<name>/channel_id     attributes digitisation, range, offset (double)
 * Group names are "read_%08u" so HDF5's name order is the input order. */
static void put_attr_f64(hid_t loc, const char *name, double v)
{
	hid_t sp = H5Screate(H5S_SCALAR);
	hid_t a = H5Acreate2(loc, name, H5T_NATIVE_DOUBLE, sp, H5P_DEFAULT, H5P_DEFAULT);
	H5Awrite(a, H5T_NATIVE_DOUBLE, &v); H5Aclose(a); H5Sclose(sp);
}
int refmap_write_fast5(const char *path, const int16_t *raw, const uint64_t *raw_off, const uint32_t *raw_len, uint32_t n_reads,
                       uint32_t first_id, double digitisation, double range, double offset)
{
	hid_t f = H5Fcreate(path, H5F_ACC_TRUNC, H5P_DEFAULT, H5P_DEFAULT);
	if (f < 0) return -1;
	for (uint32_t r = 0; r < n_reads; ++r) {
		char g[64], id[64];
		snprintf(g, sizeof(g), "/read_%08u", first_id + r);
		snprintf(id, sizeof(id), "read_%08u", first_id + r);
		hid_t gr = H5Gcreate2(f, g, H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
		hid_t raw_g = H5Gcreate2(gr, "Raw", H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
		hid_t st = H5Tcopy(H5T_C_S1); H5Tset_size(st, strlen(id) + 1);
		hid_t sp = H5Screate(H5S_SCALAR);
		hid_t a = H5Acreate2(raw_g, "read_id", st, sp, H5P_DEFAULT, H5P_DEFAULT);
		H5Awrite(a, st, id); H5Aclose(a); H5Sclose(sp); H5Tclose(st);
		hsize_t dims[1] = { raw_len[r] };
		hid_t ds_sp = H5Screate_simple(1, dims, NULL);
		hid_t ds = H5Dcreate2(raw_g, "Signal", H5T_STD_I16LE, ds_sp, H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
		H5Dwrite(ds, H5T_NATIVE_INT16, H5S_ALL, H5S_ALL, H5P_DEFAULT, raw + raw_off[r]);
		H5Dclose(ds); H5Sclose(ds_sp); H5Gclose(raw_g);
		hid_t ch = H5Gcreate2(gr, "channel_id", H5P_DEFAULT, H5P_DEFAULT, H5P_DEFAULT);
		put_attr_f64(ch, "digitisation", digitisation); put_attr_f64(ch, "range", range); put_attr_f64(ch, "offset", offset);
		H5Gclose(ch); H5Gclose(gr);
	}
	H5Fclose(f);
	return 0;
}

} /* extern "C" */
