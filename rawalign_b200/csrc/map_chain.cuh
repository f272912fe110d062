// map_chain.cuh -- chaining DP, traceback, chain selection and the PAF fields (included by map.cu only).
#pragma once

namespace mapk {

#ifndef RA_CHAIN_MIN_CTAS
#define RA_CHAIN_MIN_CTAS 8              // 60 registers, 32 warps per SM: +3.4 % on the C2 map loop (profiles/r02_ab_chain_ctas.txt)
#endif
constexpr int kMaxBest = 8;                                        // num_best_chains supported (default 3, roptions.c:18)

// traceback_chains, rmap.cpp:130-173.  score / pred / used are indexed like the keys of the segment (pred = segment index).
__device__ __forceinline__ bool traceback(const uint64_t *__restrict__ keys, const float *__restrict__ score, const uint32_t *__restrict__ pred,
                                          uint8_t *__restrict__ used, uint32_t end_idx, int min_num_anchors, ra_anchor_t *__restrict__ out,
                                          ChainRec *rec)
{
	if (used[end_idx]) return false;                               // :139
	bool stop_at_used = false;
	uint32_t start = end_idx, n = 1;
	out[0] = ra_anchor_t{ key_target(keys[start]), key_query(keys[start]) };
	if (pred[start] != start && used[pred[start]]) stop_at_used = true;
	used[start] = 1;
	while (pred[start] != start && !used[pred[start]]) {
		start = pred[start];
		out[n] = ra_anchor_t{ key_target(keys[start]), key_query(keys[start]) };
		if (pred[start] != start && used[pred[start]]) stop_at_used = true;
		used[start] = 1;
		++n;
	}
	if (n < (uint32_t)min_num_anchors) return false;               // :159
	float adj = score[end_idx];
	if (stop_at_used) adj = __fsub_rn(adj, score[pred[start]]);    // :161-163
	rec->chaining_score = adj; rec->alignment_score = 0.0f;
	rec->start_position = key_target(keys[start]); rec->end_position = key_target(keys[end_idx]);
	rec->n_anchors = n; rec->mapq = 0;
	return true;
}

__device__ __forceinline__ bool linkable(int32_t ct, int32_t cq, int32_t pt, int32_t pq, const Params &P)   // rmap.cpp:455-474
{
	const int32_t td = ct - pt, qd = cq - pq;
	if (pq == cq || pt == ct || qd < 0) return false;
	if (abs(td - qd) >= P.max_gap_length) return false;
	if (td <= 0) return true;                                      // gap_scale = 1 (:469)
	// gap_scale = RN((float)qd / td) against 0.75 and 5 (:472).  Both operands below 2^22 convert exactly and the quotient of
	// two such integers cannot lie within half an ulp of 3/4 or of 5 without being equal to it (|qd/td - 3/4| >= 1/(4 td)), so
	// the correctly rounded comparison is the integer one; larger operands take the division
	if ((uint32_t)td < (1u << 22) && (uint32_t)qd < (1u << 22)) return 4 * (int64_t)qd > 3 * (int64_t)td && (int64_t)qd < 5 * (int64_t)td;
	const float gap_scale = __fdiv_rn((float)qd, (float)td);
	return gap_scale < 5.0f && gap_scale > 0.75f;
}

// A predecessor can only be linked when 4 qd > 3 td (gap_scale > 0.75, :472) and qd = cq - pq <= cq, i.e. when 3 td < 4 cq.  The
// lists are sorted by target, so from the first predecessor with 3 td >= 4 cq on nothing can be linked any more: whatever the
// reference's scan still does there (count skips until the target-gap break or the cut-off) cannot change the anchor's score or
// predecessor.  The scans below stop there -- for the anchors of a first chunk (cq < ~500 events) a window of < 670 target
// positions instead of max_target_gap_length = 5000.
__device__ __forceinline__ bool beyond_link_reach(int32_t ct, int32_t cq, int32_t pt) { return 3ll * ((long long)ct - pt) >= 4ll * cq; }

// Pre-pass of the chaining DP, one lane per anchor.  Most anchors are isolated seed hits: no predecessor the reference's scan
// would look at (:452-484: up to the target-gap break, the chaining band, or max_num_skips + 1 comparisons that cannot
// improve) satisfies the gap conditions of :472, so their score stays the seed value e and they are their own predecessor --
// decided here, in parallel, without the recurrence.  The anchors that do have a linkable predecessor in reach are listed in
// order (nt) for the sequential pass.  Every anchor gets score = e, pred = self, used = 0.
constexpr int kPreThreads = 256;
__global__ void __launch_bounds__(kPreThreads)
k_chain_pre(const uint32_t *__restrict__ act, uint32_t n_act, Params P, const uint64_t *__restrict__ keys, const uint64_t *__restrict__ seg_base,
            const uint32_t *__restrict__ seg_n, const uint32_t *__restrict__ skip, float *__restrict__ score, uint32_t *__restrict__ pred,
            uint8_t *__restrict__ used, uint32_t *__restrict__ nt, uint32_t *__restrict__ n_nt, uint8_t *__restrict__ ntf)
{
	__shared__ uint32_t warp_cnt[kPreThreads / 32];
	const uint32_t i = blockIdx.x;                                 // one CTA per read: 8 warps keep 8 scans in flight per read
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	if (i >= n_act) return;
	if (skip[i]) { if (tid == 0) n_nt[i] = 0; return; }
	const uint64_t base = seg_base[i];
	const uint32_t n = seg_n[i];
	const uint64_t *__restrict__ K = keys + base;
	const float fe = (float)P.e;
	const bool all = P.no_filter || fe >= P.min_chaining_score;    // an isolated anchor could itself be a chain end: no shortcut
	uint32_t cnt = 0;
	for (uint32_t a0 = 0; a0 < n; a0 += kPreThreads) {
		const uint32_t a = a0 + tid;
		const bool mine = a < n;
		const uint64_t kc = mine ? K[a] : 0ull;
		const uint32_t list = key_list(kc);
		const int32_t ct = (int32_t)key_target(kc), cq = (int32_t)key_query(kc);
		if (mine) { score[base + a] = fe; pred[base + a] = a; used[base + a] = 0; }
		bool hit = mine && all;
		bool active = mine && !all && a > 0;
		uint32_t pi = a;
		int32_t looked = 0;
		// the lanes walk their predecessors in lock-step (a warp-uniform loop with predicated bodies: per-lane break / continue
		// would leave the lanes diverged, 32 times the latency)
		while (__any_sync(0xffffffffu, active)) {
			const uint32_t pj = active ? pi - 1u : 0u;
			const uint64_t kp = K[pj];
			const int32_t pt = (int32_t)key_target(kp), pq = (int32_t)key_query(kp);
			const bool same = key_list(kp) == list && a - pj <= (uint32_t)P.chaining_band_length;
			const bool skipc = pq == cq || pt == ct;                                          // :455-456
			const bool brk = (!skipc && (pt + P.max_target_gap_length < ct)) || beyond_link_reach(ct, cq, pt);   // :458
			const bool consider = active && same && !skipc && !brk && (cq - pq) >= 0;         // :465
			const bool link = consider && linkable(ct, cq, pt, pq, P);
			hit |= link;
			looked += (consider && !link) ? 1 : 0;
			active = active && same && !brk && !link && looked <= P.max_num_skips && pj > 0;  // :482: the reference's scan has given up by then
			pi = pj;
		}
		if (mine) ntf[base + a] = hit ? 1 : 0;
		const uint32_t m = __ballot_sync(0xffffffffu, hit);        // ordered compaction of the tile's listed anchors
		if (lane == 0) warp_cnt[wid] = __popc(m);
		__syncthreads();
		uint32_t before = 0, total = 0;
#pragma unroll
		for (int w = 0; w < kPreThreads / 32; ++w) { const uint32_t t = warp_cnt[w]; if (w < wid) before += t; total += t; }
		if (hit) nt[base + cnt + before + __popc(m & ((1u << lane) - 1u))] = a;
		cnt += total;
		__syncthreads();
	}
	if (tid == 0) n_nt[i] = cnt;
}

// Second parallel pass, one lane per LISTED anchor.  Most listed anchors are chance pairs: every linkable predecessor the
// reference's scan reaches is itself an isolated anchor, whose score is known to be e.  For those the whole scan of
// :452-484 -- strict-improvement rule, skip counter, cut-off -- is replayed here by the lane on its own (lock-step
// predicated loop), and score / predecessor are final.  An anchor whose scan touches a linkable predecessor that is itself
// listed depends on a score not known yet: it is marked complex (cx = 1) and left to the sequential pass.
__global__ void __launch_bounds__(kPreThreads)
k_chain_simple(const uint32_t *__restrict__ act, uint32_t n_act, Params P, const uint64_t *__restrict__ keys, const uint64_t *__restrict__ seg_base,
               const uint32_t *__restrict__ skip, float *__restrict__ score, uint32_t *__restrict__ pred, const uint32_t *__restrict__ nt,
               const uint32_t *__restrict__ n_nt, const uint8_t *__restrict__ ntf, uint8_t *__restrict__ cx)
{
	const uint32_t i = blockIdx.x;
	const int tid = threadIdx.x;
	if (i >= n_act || skip[i]) return;
	const uint64_t base = seg_base[i];
	const uint64_t *__restrict__ K = keys + base;
	const uint8_t *__restrict__ F = ntf + base;
	const uint32_t m_nt = n_nt[i];
	const float fe = (float)P.e;
	for (uint32_t j0 = 0; j0 < m_nt; j0 += kPreThreads) {
		const uint32_t j = j0 + tid;
		const bool mine = j < m_nt;
		const uint32_t a = mine ? nt[base + j] : 0u;
		const uint64_t kc = K[a];
		const uint32_t list = key_list(kc);
		const int32_t ct = (int32_t)key_target(kc), cq = (int32_t)key_query(kc);
		float sc = fe; uint32_t pr = a;
		int32_t skips = 0;
		bool complex_ = false;
		bool active = mine && a > 0;
		uint32_t pi = a;
		while (__any_sync(0xffffffffu, active)) {
			const uint32_t pj = active ? pi - 1u : 0u;
			const uint64_t kp = K[pj];
			const int32_t pt = (int32_t)key_target(kp), pq = (int32_t)key_query(kp);
			const bool same = key_list(kp) == list && a - pj <= (uint32_t)P.chaining_band_length;    // :446-448
			const bool skipc = pq == cq || pt == ct;                                          // :455-456
			const bool brk = (!skipc && (pt + P.max_target_gap_length < ct)) || beyond_link_reach(ct, cq, pt);   // :458
			const bool consider = active && same && !skipc && !brk && (cq - pq) >= 0;         // :465
			const bool link = consider && linkable(ct, cq, pt, pq, P);
			if (link && F[pj]) complex_ = true;                                                // its score is not e and not known yet
			const float cur = link ? __fadd_rn(fe, (float)min(min(ct - pt, cq - pq), P.e)) : 0.0f;   // :467-474 with score[pj] = e
			const bool better = consider && cur > sc;                                          // :476
			if (better) { sc = cur; pr = pj; }
			skips += consider ? (better ? -1 : 1) : 0;                                         // :479-481
			active = active && same && !brk && !complex_ && !(consider && !better && skips > P.max_num_skips) && pj > 0;   // :482
			pi = pj;
		}
		if (mine) {
			cx[base + j] = complex_ ? 1 : 0;
			if (!complex_) { score[base + a] = sc; pred[base + a] = pr; }
		}
	}
}

// Chaining DP of gen_chains (rmap.cpp:427-507) for every (ref, strand) list of one read, in the reference's list order,
// with the running max_chaining_score shared by the lists (:428).  One warp per read, over the anchors k_chain_pre listed.
// The anchors of a list are walked in order (score[ai] needs the scores before it), but the predecessor scan of one anchor
// (:452-484) is done 32 predecessors at a time: lane l looks at anchor ai-1-l (one coalesced load), and the loop's sequential
// semantics are rebuilt from warp prefix operations --
//   * the target-gap break (:458) is monotonic along the lanes (the list is sorted by target): the first lane that hits it,
//     or that leaves the chaining band / the list, ends the scan;
//   * score[ai] only ever grows, so a predecessor improves it iff its candidate score beats the seed value and every
//     candidate before it: an exclusive prefix maximum;
//   * the skip counter (:479-482) is a prefix sum of -1 (improved) / +1 (did not) over the lanes that reach the comparison;
//     the first non-improving lane that takes it past max_num_skips is the last one the reference looks at.
// All lanes carry the same running state (score, predecessor, skip count, best-candidate list), lane 0 writes and does the
// tracebacks.  The running maximum only needs the listed anchors: every score is >= e, and the filter of :488-492 is applied
// after the anchor's own score has entered the maximum.  Candidate chains go to cand[i * cap + k] in generation order, or
// sorted by chaining score (stable, :512) when they are to be evaluated by DTW.
__global__ void __launch_bounds__(128, RA_CHAIN_MIN_CTAS)
k_chain(const uint32_t *__restrict__ act, uint32_t n_act, Params P, const uint64_t *__restrict__ keys, const uint64_t *__restrict__ seg_base,
        const uint32_t *__restrict__ seg_n, const uint32_t *__restrict__ skip, float *__restrict__ score, uint32_t *__restrict__ pred,
        uint8_t *__restrict__ used, const uint32_t *__restrict__ nt, const uint32_t *__restrict__ n_nt, const uint8_t *__restrict__ cx,
        ra_anchor_t *__restrict__ anc_cur, ChainRec *__restrict__ cand, uint32_t *__restrict__ n_cand)
{
	const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (i >= n_act) return;
	if (skip[i]) { if (lane == 0) n_cand[i] = 0; return; }
	const uint64_t base = seg_base[i];
	const uint64_t *__restrict__ K = keys + base;
	float *S = score + base;
	uint32_t *__restrict__ Pd = pred + base;
	uint8_t *__restrict__ U = used + base;
	const uint32_t *__restrict__ NT = nt + base;
	const uint8_t *__restrict__ CX = cx + base;
	const uint32_t m_nt = n_nt[i];
	ChainRec *out = cand + (size_t)i * P.cap_chains;
	uint32_t nc = 0;                                               // lane 0 only
	uint64_t wo = base;                                            // next free chain-anchor slot of this read (lane 0 only)
	float max_chaining_score = 0.0f;                               // :428
	const float fe = (float)P.e;
	// the listed anchors are taken 32 at a time: lane l holds entry jb + l (index, key, complex flag, and -- for the anchors
	// k_chain_simple finished -- the final score); the walk below reads them with shuffles
	uint32_t jb = 0, nt_reg = 0, cx_reg = 0; float s_reg = 0.0f; uint64_t k_reg = 0;
	auto load_block = [&](uint32_t j0) {
		jb = j0;
		const bool in = jb + lane < m_nt;
		nt_reg = in ? NT[jb + lane] : 0u; cx_reg = in ? (uint32_t)CX[jb + lane] : 0u;
		k_reg = in ? K[nt_reg] : 0ull; s_reg = in ? S[nt_reg] : 0.0f;
	};
	if (m_nt) load_block(0);
	uint32_t j = 0;
	while (j < m_nt) {
		if (j - jb >= 32u) load_block(j);
		const uint32_t first = __shfl_sync(0xffffffffu, nt_reg, (int)(j - jb));
		const uint32_t list = key_list(__shfl_sync(0xffffffffu, k_reg, (int)(j - jb)));
		uint32_t ls;                                               // first anchor of this (ref, strand) list: lower bound of the list id
		{
			uint32_t lo = 0, hi = first;
			while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (key_list(K[mid]) < list) lo = mid + 1; else hi = mid; }
			ls = lo;
		}
		float cs[kMaxBest]; uint32_t ci[kMaxBest]; int ncand = 0;
#pragma unroll
		for (int t = 0; t < kMaxBest; ++t) { cs[t] = 0.0f; ci[t] = 0u; }
		// inserts (sc, ai) into the num_best_chains best by (score desc, index desc) (compare(), :175-179): ai only grows, so a tie goes
		// in front; fully unrolled with selects: the list stays in registers
		auto keep_best = [&](float sc, uint32_t ai) {
			int p = 0;
#pragma unroll
			for (int t = 0; t < kMaxBest; ++t) p += (t < ncand && cs[t] > sc) ? 1 : 0;
			if (p < P.num_best_chains) {
				const int lastp = min(ncand, P.num_best_chains - 1);
#pragma unroll
				for (int t = kMaxBest - 1; t > 0; --t) { const bool sh = t > p && t <= lastp; cs[t] = sh ? cs[t - 1] : cs[t]; ci[t] = sh ? ci[t - 1] : ci[t]; }
#pragma unroll
				for (int t = 0; t < kMaxBest; ++t) { const bool here = t == p; cs[t] = here ? sc : cs[t]; ci[t] = here ? ai : ci[t]; }
				if (ncand < P.num_best_chains) ++ncand;
			}
		};
		while (j < m_nt) {
			if (j - jb >= 32u) load_block(j);
			const int sl = (int)(j - jb);
			// a run of listed anchors of this list that k_chain_simple finished is taken in one step: their scores are final, the
			// running maximum after each of them is a prefix maximum, and the filter of :488-492 lets through a handful at most
			{
				const bool in_list = jb + (uint32_t)lane < m_nt && key_list(k_reg) == list;
				const uint32_t stop = __ballot_sync(0xffffffffu, !in_list || cx_reg != 0u) & ~((1u << sl) - 1u);
				const int run_end = stop ? __ffs(stop) - 1 : 32;
				if (run_end > sl) {
					const bool in_run = lane >= sl && lane < run_end;
					float pm = in_run ? s_reg : 0.0f;                  // scores are >= e > 0
#pragma unroll
					for (int d = 1; d < 32; d <<= 1) { const float y = __shfl_up_sync(0xffffffffu, pm, d); if (lane >= d) pm = fmaxf(pm, y); }
					pm = fmaxf(pm, max_chaining_score);                // the maximum once this lane's score has entered it (:485-487)
					const bool pass = in_run && (P.no_filter || (s_reg >= P.min_chaining_score && s_reg > pm / 2));   // :488-492
					uint32_t pass_mask = __ballot_sync(0xffffffffu, pass);
					max_chaining_score = __shfl_sync(0xffffffffu, pm, run_end - 1);
					while (pass_mask) {
						const int t = __ffs(pass_mask) - 1; pass_mask &= pass_mask - 1;
						keep_best(__shfl_sync(0xffffffffu, s_reg, t), __shfl_sync(0xffffffffu, nt_reg, t));
					}
					j += (uint32_t)(run_end - sl);
					continue;
				}
			}
			const uint32_t ai = __shfl_sync(0xffffffffu, nt_reg, sl);
			const uint64_t kc = __shfl_sync(0xffffffffu, k_reg, sl);
			if (key_list(kc) != list) break;
			float sc;
			{                                                      // a complex anchor: the scan of :452-484, 32 predecessors at a time
				sc = fe; uint32_t pr = ai;                         // :442-443
				const int32_t ct = (int32_t)key_target(kc), cq = (int32_t)key_query(kc);
				uint32_t start = ls;
				if (ai - ls > (uint32_t)P.chaining_band_length) start = ai - (uint32_t)P.chaining_band_length;   // :448
				int32_t num_skips = 0;
				bool done = false;
				for (uint32_t top = ai; top > start && !done; top = top >= 32 ? top - 32 : 0) {
					const bool inrange = top >= start + 1u + (uint32_t)lane;          // predecessor top-1-lane is still >= start
					const uint32_t pi = inrange ? top - 1u - (uint32_t)lane : start;
					const uint64_t kp = K[pi];
					const int32_t pt = (int32_t)key_target(kp), pq = (int32_t)key_query(kp);
					const bool brk_t = inrange && ((pt + P.max_target_gap_length < ct) || beyond_link_reach(ct, cq, pt));   // :458
					const uint32_t dead_mask = __ballot_sync(0xffffffffu, brk_t || !inrange);
					const int first_dead = dead_mask ? __ffs(dead_mask) - 1 : 32;
					const int32_t td = ct - pt, qd = cq - pq;
					const bool valid = lane < first_dead && pq != cq && pt != ct && qd >= 0;     // :455-457, :465
					float cur = 0.0f;
					if (valid && linkable(ct, cq, pt, pq, P)) cur = __fadd_rn(S[pi], (float)min(min(td, qd), P.e));   // :467-474 (stored by an earlier kernel, or by lane 0 before a __syncwarp)
					const uint32_t valid_mask = __ballot_sync(0xffffffffu, valid);
					const uint32_t beat_mask = __ballot_sync(0xffffffffu, cur > sc);
					if (beat_mask != 0 && (beat_mask & (beat_mask - 1)) == 0) {
						// exactly one lane t beats the score so far: it is the only improvement (-1), every other counting lane a skip (+1)
						const int t = __ffs(beat_mask) - 1;
						const int room = P.max_num_skips + 1 - num_skips;
						const int before_t = __popc(valid_mask & ((1u << t) - 1u));
						if (before_t >= room) { done = true; continue; }              // the scan ends before it reaches lane t (:482)
						sc = __shfl_sync(0xffffffffu, cur, t); pr = __shfl_sync(0xffffffffu, pi, t);
						num_skips += before_t - 1;
						const int after_t = __popc(valid_mask >> t >> 1);
						if (after_t >= P.max_num_skips + 1 - num_skips) done = true; else num_skips += after_t;
						if (first_dead < 32) done = true;
						continue;
					}
					if (beat_mask == 0) {
						// nothing here beats the score so far: every lane that counts is a skip (:479-482)
						const int room = P.max_num_skips + 1 - num_skips;             // the room-th skip from here ends the scan
						const int have = __popc(valid_mask);
						if (have >= room) done = true; else num_skips += have;
						if (first_dead < 32) done = true;
						continue;
					}
					float pm = cur;                                    // inclusive prefix maximum (0 for the lanes that do not count)
#pragma unroll
					for (int d = 1; d < 32; d <<= 1) { const float y = __shfl_up_sync(0xffffffffu, pm, d); if (lane >= d) pm = fmaxf(pm, y); }
					float before = __shfl_up_sync(0xffffffffu, pm, 1);
					before = fmaxf(lane ? before : 0.0f, sc);
					const bool improve = valid && cur > before;        // :476
					int32_t ps = valid ? (improve ? -1 : 1) : 0;       // :479-481
#pragma unroll
					for (int d = 1; d < 32; d <<= 1) { const int32_t y = __shfl_up_sync(0xffffffffu, ps, d); if (lane >= d) ps += y; }
					ps += num_skips;
					const uint32_t brk_mask = __ballot_sync(0xffffffffu, valid && !improve && ps > P.max_num_skips);   // :482
					const int last = brk_mask ? __ffs(brk_mask) - 1 : 31;
					const uint32_t imp_mask = __ballot_sync(0xffffffffu, improve && lane <= last);
					if (imp_mask) {
						const int t = 31 - __clz(imp_mask);            // improvements grow along the lanes: the last one holds the maximum
						sc = __shfl_sync(0xffffffffu, cur, t); pr = __shfl_sync(0xffffffffu, pi, t);
					}
					num_skips = __shfl_sync(0xffffffffu, ps, last);
					done = brk_mask != 0 || first_dead < 32;
				}
				if (lane == 0) { S[ai] = sc; Pd[ai] = pr; }
				__syncwarp();
			}
			if (sc > max_chaining_score) max_chaining_score = sc;  // :485-487
			if (P.no_filter || (sc >= P.min_chaining_score && sc > max_chaining_score / 2)) keep_best(sc, ai);   // :488-492
			++j;
		}
		if (lane == 0) {
			bool stop = false;
#pragma unroll
			for (int k = 0; k < kMaxBest; ++k) {                   // :498-505
				if (k >= ncand || stop) continue;
				ChainRec rec;
				if (nc < P.cap_chains && traceback(K, S, Pd, U, ci[k], P.min_num_anchors, anc_cur + wo, &rec)) {
					rec.ref = list >> 1; rec.strand = list & 1u; rec.anchor_off = wo;
					wo += rec.n_anchors;
					out[nc++] = rec;
				}
				if (!P.no_filter && S[ci[k]] < max_chaining_score / 2) stop = true;
			}
		}
		__syncwarp();
	}
	if (lane != 0) return;
	if (P.dtw)                                                     // :512, insertion sort = libstdc++'s std::sort for <= 16 elements
		for (uint32_t a = 1; a < nc; ++a) {
			const ChainRec x = out[a];
			uint32_t b = a;
			while (b > 0 && x.chaining_score > out[b - 1].chaining_score) { out[b] = out[b - 1]; --b; }
			out[b] = x;
		}
	n_cand[i] = nc;
}

// the candidates as ra_chain_t for ra_align_chains_batch_dev: one group per active read (possibly empty)
__global__ void __launch_bounds__(128)
k_emit_dtw_chains(const ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, Params P, const ChainRec *__restrict__ cand,
                  const uint32_t *__restrict__ n_cand, const uint64_t *__restrict__ cand_base, const uint64_t *__restrict__ seq_tab,
                  ra_chain_t *__restrict__ out, uint32_t *__restrict__ group_first)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i > n_act) return;
	group_first[i] = (uint32_t)cand_base[i];
	if (i == n_act) return;
	const uint32_t r = act[i];
	const ChainRec *c = cand + (size_t)i * P.cap_chains;
	const uint64_t b = cand_base[i];
	for (uint32_t k = 0; k < n_cand[i]; ++k) {
		ra_chain_t o;
		o.anchor_off = c[k].anchor_off; o.n_anchors = c[k].n_anchors; o.group = i; o.read_off = st[r].ev_base;
		o.ref_off = seq_tab[3 * (size_t)c[k].ref + (c[k].strand == 1 ? 0 : 1)];                     // rmap.cpp:183-188
		out[b + k] = o;
	}
}

__device__ __forceinline__ bool tuple_greater(const ChainRec &a, const ChainRec &b)                  // rmap.h:41-45
{
	if (a.alignment_score != b.alignment_score) return a.alignment_score > b.alignment_score;
	if (a.chaining_score != b.chaining_score) return a.chaining_score > b.chaining_score;
	if (a.n_anchors != b.n_anchors) return a.n_anchors > b.n_anchors;
	if (a.strand != b.strand) return a.strand > b.strand;
	if (a.ref != b.ref) return a.ref > b.ref;
	if (a.start_position != b.start_position) return a.start_position > b.start_position;
	return a.end_position > b.end_position;
}

// is_mapped_with_high_confidence, rmap.cpp:594-665
__device__ __forceinline__ bool is_mapped(const ChainRec *c, uint32_t n_chains, const Params &P)
{
	const uint32_t n_anchors0 = n_chains ? c[0].n_anchors : 0u;
	if (n_anchors0 == 0) return false;
#define RA_SC(x) (P.dtw ? (x).alignment_score : (x).chaining_score)
	if (n_chains >= 2) {
		if (__fdiv_rn(RA_SC(c[0]), RA_SC(c[1])) >= P.min_bestmap_ratio) return true;
		float mean = 0.0f;
		for (uint32_t k = 0; k < n_chains; ++k) mean = __fadd_rn(mean, RA_SC(c[k]));
		mean = __fdiv_rn(mean, (float)n_chains);
		if (RA_SC(c[0]) >= __fmul_rn(P.min_meanmap_ratio, mean)) return true;
	} else if (n_chains == 1 && c[0].n_anchors >= P.min_chain_anchor) return true;
#undef RA_SC
	return false;
}

// After the DTW evaluation of a round: the filter of rmap.cpp:518-523, gen_primary_chains (:90-128), comp_mapq (:65-88), the
// loop control of :685-693 and, for reads leaving the loop, the PAF fields of :696-800.  One thread per active read.
__global__ void __launch_bounds__(64)
k_select(ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, Params P, ChainRec *__restrict__ cand,
         const uint32_t *__restrict__ n_cand, const uint64_t *__restrict__ cand_base, const float *__restrict__ dscore, const uint8_t *__restrict__ dkeep,
         const uint32_t *__restrict__ skip, ChainRec *__restrict__ chains, const ra_anchor_t *__restrict__ anc_cur, uint32_t cur_arena,
         const uint64_t *__restrict__ seq_tab, ra_paf_t *__restrict__ paf, uint32_t *__restrict__ counters /* [2] still active */)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_act) return;
	const uint32_t r = act[i];
	ReadState s = st[r];
	ChainRec *rc = chains + (size_t)r * P.cap_chains;
	if (!skip[i]) {
		ChainRec *c = cand + (size_t)i * P.cap_chains;
		uint32_t n = n_cand[i];
		if (P.dtw) {                                               // :515-527
			const uint64_t b = cand_base[i];
			uint32_t m = 0;
			for (uint32_t k = 0; k < n; ++k) {
				ChainRec x = c[k];
				x.alignment_score = dscore[b + k];
				if (dkeep[b + k]) c[m++] = x;
			}
			n = m;
		}
		uint32_t np = 0;
		if (n > 0) {
			for (uint32_t a = 1; a < n; ++a) {                     // std::sort(greater) :91 (stable insertion)
				const ChainRec x = c[a];
				uint32_t b = a;
				while (b > 0 && tuple_greater(x, c[b - 1])) { c[b] = c[b - 1]; --b; }
				c[b] = x;
			}
			rc[np++] = c[0];
			for (uint32_t ci = 1; ci < n; ++ci) {
				const ChainRec &last = rc[np - 1];
				if (P.dtw) { if (c[ci].alignment_score < __fdiv_rn(last.alignment_score, 3.0f)) break; }   // :100-104
				else if (c[ci].chaining_score < __fdiv_rn(last.chaining_score, 3.0f)) break;
				bool primary = true;
				for (uint32_t pi = 0; pi < np; ++pi)
					if (c[ci].ref == rc[pi].ref && max(c[ci].start_position, rc[pi].start_position) <= min(c[ci].end_position, rc[pi].end_position)) { primary = false; break; }
				if (primary) rc[np++] = c[ci];
			}
			if (np == 1) rc[0].mapq = 60;                          // comp_mapq
			else {
				const float q = P.dtw ? __fdiv_rn(rc[1].alignment_score, rc[0].alignment_score) : __fdiv_rn(rc[1].chaining_score, rc[0].chaining_score);
				int mapq = (int)__fmul_rn(40.0f, __fsub_rn(1.0f, q));
				mapq = min(60, max(0, mapq));
				rc[0].mapq = (uint32_t)mapq;
			}
		}
		s.n_chains = np; s.arena = cur_arena;
	}
	const bool mapped = is_mapped(rc, s.n_chains, P);
	bool leave = mapped;
	if (!mapped) {                                                 // :685: chunk_start += l_chunk, ++current_chunk
		s.chunk_start += P.chunk_size; s.current_chunk += 1;
		if (s.chunk_start >= s.l_sig || s.current_chunk >= P.max_num_chunk) leave = true;
	}
	if (leave) {
		s.active = 0; s.done = 1;
		uint32_t current_chunk = s.current_chunk;
		if (current_chunk > 0 && (s.chunk_start >= s.l_sig || current_chunk == P.max_num_chunk)) --current_chunk;   // :696
		const float scale = __fdiv_rn(__fdiv_rn(__fmul_rn((float)(current_chunk + 1), (float)P.chunk_size), (float)s.offset),
		                              __fdiv_rn((float)P.sample_rate, (float)P.bp_per_sec));                      // :698
		const uint32_t n_chains = s.n_chains;
		const uint32_t n_anchors0 = n_chains ? rc[0].n_anchors : 0u;
		float mean_chain_score = 0.0f;
		for (uint32_t k = 0; k < n_chains; ++k) mean_chain_score = __fadd_rn(mean_chain_score, rc[k].chaining_score);
		mean_chain_score = __fdiv_rn(mean_chain_score, (float)n_chains);
		ra_paf_t o; memset(&o, 0, sizeof(o));
		o.ci = current_chunk + 1; o.sl = s.l_sig; o.n_events = s.n_events;
		float at = 0.0f, aq = 0.0f;
		const ra_anchor_t *a0 = n_chains ? anc_cur + rc[0].anchor_off : nullptr;
		for (uint32_t ai = 0; ai + 1 < n_anchors0; ++ai) {
			at = __fadd_rn(at, (float)(a0[ai].target_position - a0[ai + 1].target_position));
			aq = __fadd_rn(aq, (float)(a0[ai].query_position - a0[ai + 1].query_position));
		}
		if (mapped) {
			at = __fdiv_rn(at, (float)n_anchors0); aq = __fdiv_rn(aq, (float)n_anchors0);
			o.mapped = 1; o.cm = n_anchors0; o.nc = n_chains; o.s1 = rc[0].chaining_score; o.s2 = n_chains > 1 ? rc[1].chaining_score : 0.0f;
			o.sm = mean_chain_score; o.at = at; o.aq = aq; o.alns = rc[0].alignment_score;
			o.ref_id = rc[0].ref;
			o.read_length = (uint32_t)__fmul_rn(scale, (float)a0[0].query_position);
			o.read_start_position = (uint32_t)__fmul_rn(scale, (float)a0[n_anchors0 - 1].query_position);
			o.read_end_position = (uint32_t)__fmul_rn(scale, (float)a0[0].query_position);
			const uint32_t ref_len = (uint32_t)seq_tab[3 * (size_t)rc[0].ref + 2];
			o.fragment_start_position = rc[0].strand ? (ref_len + 1u - rc[0].end_position) : rc[0].start_position;
			o.fragment_length = rc[0].end_position - rc[0].start_position + 1u;
			o.mapq = rc[0].mapq & 63u; o.rev = rc[0].strand & 1u;
		} else {
			if (n_chains >= 1) {
				if (n_anchors0) { at = __fdiv_rn(at, (float)n_anchors0); aq = __fdiv_rn(aq, (float)n_anchors0); }
				o.cm = n_anchors0; o.nc = n_chains; o.s1 = rc[0].chaining_score; o.s2 = n_chains > 1 ? rc[1].chaining_score : 0.0f;
				o.sm = mean_chain_score; o.at = at; o.aq = aq;
			}
			o.read_length = (uint32_t)__fmul_rn(scale, (float)s.offset);
		}
		paf[r] = o;
	} else atomicAdd(&counters[2], 1u);
	st[r] = s;
}

// reg->offset += n_chunk_events for the reads that chained this round (rmap.cpp:575), before k_select looks at it
__global__ void k_advance_offset(ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, const uint64_t *__restrict__ evoff,
                                 const uint32_t *__restrict__ skip)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_act || skip[i]) return;
	st[act[i]].offset += (uint32_t)(evoff[i + 1] - evoff[i]);
}

// ---- --dtw-output-cigar (rmap.cpp:715-717): chains[0] of the reads that left the loop mapped in this round -------------
// ordered list of those reads + exclusive scan of their chain-0 anchor counts; one CTA
__global__ void __launch_bounds__(1024)
k_cigar_list(const ReadState *__restrict__ st, const uint32_t *__restrict__ act, uint32_t n_act, Params P, const ChainRec *__restrict__ chains,
             const ra_paf_t *__restrict__ paf, uint32_t *__restrict__ fin_r, uint64_t *__restrict__ fin_aoff, uint32_t *__restrict__ counters /* [4] n_fin, [5] anchors */)
{
	__shared__ uint32_t wc[32], wa[32];
	__shared__ uint32_t base_c, base_a;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	if (tid == 0) { base_c = 0; base_a = 0; }
	__syncthreads();
	for (uint32_t i0 = 0; i0 < n_act; i0 += 1024) {
		const uint32_t i = i0 + tid;
		uint32_t r = 0, na = 0; bool fin = false;
		if (i < n_act) { r = act[i]; fin = !st[r].active && paf[r].mapped; if (fin) na = chains[(size_t)r * P.cap_chains].n_anchors; }
		const uint32_t m = __ballot_sync(0xffffffffu, fin);
		uint32_t inc = na;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += y; }
		if (lane == 31) wa[wid] = inc;
		if (lane == 0) wc[wid] = __popc(m);
		__syncthreads();
		uint32_t bc = 0, ba = 0, tc = 0, ta = 0;
		for (int w = 0; w < 32; ++w) { if (w < wid) { bc += wc[w]; ba += wa[w]; } tc += wc[w]; ta += wa[w]; }
		if (fin) {
			const uint32_t k = base_c + bc + __popc(m & ((1u << lane) - 1u));
			fin_r[k] = r; fin_aoff[k] = base_a + ba + inc - na;
		}
		__syncthreads();
		if (tid == 0) { base_c += tc; base_a += ta; }
		__syncthreads();
	}
	if (tid == 0) { counters[4] = base_c; counters[5] = base_a; }
}

// the listed chains as ra_chain_t + their anchors, contiguous; one warp per chain
__global__ void __launch_bounds__(128)
k_cigar_pack(const ReadState *__restrict__ st, Params P, const ChainRec *__restrict__ chains, const ra_anchor_t *__restrict__ anc_cur,
             const uint64_t *__restrict__ seq_tab, const uint32_t *__restrict__ fin_r, const uint64_t *__restrict__ fin_aoff, uint32_t n_fin,
             ra_chain_t *__restrict__ out_chain, ra_anchor_t *__restrict__ out_anc)
{
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (k >= n_fin) return;
	const uint32_t r = fin_r[k];
	const ChainRec c = chains[(size_t)r * P.cap_chains];
	const uint64_t o = fin_aoff[k];
	for (uint32_t a = lane; a < c.n_anchors; a += 32) out_anc[o + a] = anc_cur[c.anchor_off + a];
	if (lane == 0) {
		ra_chain_t ch;
		ch.anchor_off = o; ch.n_anchors = c.n_anchors; ch.group = k; ch.read_off = st[r].ev_base;
		ch.ref_off = seq_tab[3 * (size_t)c.ref + (c.strand == 1 ? 0 : 1)];
		out_chain[k] = ch;
	}
}

} // namespace mapk
