// map_batch.cuh -- host orchestration of ra_map_batch (included by map.cu only).
#pragma once
#include <chrono>

namespace mapk {

__global__ void k_init_state(ReadState *__restrict__ st, uint32_t n_reads, const uint32_t *__restrict__ l_sig_dev, ra_paf_t *__restrict__ paf)
{
	const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= n_reads) return;
	ReadState s = st[r];
	if (l_sig_dev) s.l_sig = l_sig_dev[r];
	s.active = s.l_sig > 0; s.done = s.l_sig == 0;
	ra_paf_t o; memset(&o, 0, sizeof(o));
	if (s.l_sig == 0) { o.ci = 1; o.sl = 0; }                      // the loop of rmap.cpp:685 does not run: an unmapped line with ci:i:1
	paf[r] = o;
	st[r] = s;
}

// Float-signal form, per-round upload: the round's chunk of every active read is fetched straight from the caller's buffer when
// that is page-locked (mapped host memory read over PCIe, 128-byte coalesced) into a dense round arena, chunk i at i * stride.
__global__ void __launch_bounds__(256)
k_gather_chunks(const float *__restrict__ hsig, uint64_t *__restrict__ coff, const uint32_t *__restrict__ clen, uint32_t n_act, float *__restrict__ dst, uint32_t stride)
{
	// a small persistent grid (PCIe needs ~100 KB in flight, not the whole machine: the other batches' kernels keep their SM slots),
	// one warp per chunk at a time, four 16-byte loads in flight per lane
	const int lane = threadIdx.x & 31;
	const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t i = warp; i < n_act; i += n_warps) {
		const float *__restrict__ src = hsig + coff[i];
		const uint32_t n = clen[i];
		float *__restrict__ d = dst + (size_t)i * stride;
		const uint32_t head = min(n, (uint32_t)(((16u - (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u)) & 15u) >> 2));
		if ((uint32_t)lane < head) d[lane] = src[lane];
		const float4 *__restrict__ s4 = reinterpret_cast<const float4 *>(src + head);
		const uint32_t n4 = (n - head) >> 2;
		uint32_t k = lane;
		for (; k + 96 < n4; k += 128) {
			const float4 v0 = s4[k], v1 = s4[k + 32], v2 = s4[k + 64], v3 = s4[k + 96];
			float *o = d + head + 4u * k;
			o[0] = v0.x; o[1] = v0.y; o[2] = v0.z; o[3] = v0.w;
			o[128] = v1.x; o[129] = v1.y; o[130] = v1.z; o[131] = v1.w;
			o[256] = v2.x; o[257] = v2.y; o[258] = v2.z; o[259] = v2.w;
			o[384] = v3.x; o[385] = v3.y; o[386] = v3.z; o[387] = v3.w;
		}
		for (; k < n4; k += 32) {
			const float4 v = s4[k];
			float *o = d + head + 4u * k;
			o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
		}
		const uint32_t tail0 = head + 4u * n4;
		if (tail0 + (uint32_t)lane < n) d[tail0 + lane] = src[tail0 + lane];
		__syncwarp();
		if (lane == 0) coff[i] = (uint64_t)i * stride;
	}
}

__global__ void k_dense_coff(uint64_t *__restrict__ coff, uint32_t n_act, uint32_t stride)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n_act) coff[i] = (uint64_t)i * stride;
}

} // namespace mapk

namespace {

struct StageClock {
	std::chrono::steady_clock::time_point t;
	void start() { t = std::chrono::steady_clock::now(); }
	float lap() { auto n = std::chrono::steady_clock::now(); float ms = std::chrono::duration<float, std::milli>(n - t).count(); t = n; return ms; }
};

inline uint32_t events_bound(uint32_t len, uint32_t w1, uint32_t w2) { return len / (w1 / 2 + 2) + len / (w2 / 2 + 2) + 4; }

} // namespace

extern "C" int ra_map_batch(ra_ctx *ctx, ra_index *ix, const ra_ri_mapopt_t *opt,
                            const int16_t *raw, const uint64_t *raw_off, const uint32_t *raw_len, const float *offset, const float *scale,
                            const float *sig, const uint64_t *sig_off, const uint32_t *sig_len,
                            uint32_t n_reads, ra_paf_t *out, ra_map_stats_t *stats)
{
	using namespace mapk;
	if (!ctx || !ix || !opt || !out) return ra__fail(RA_EINVAL, "NULL argument");
	if (ix->ctx != ctx) return ra__fail(RA_EINVAL, "the index was loaded for another context");
	const bool from_raw = raw != nullptr;
	const bool resident = !raw && !sig;                            // map again the signal the previous call left on the device
	if (resident) {
		if (ix->last_len.size() != n_reads || n_reads == 0) return ra__fail(RA_EINVAL, "no resident batch of %u reads on the device", n_reads);
	} else if (from_raw ? (!raw_off || !raw_len || !offset || !scale) : (!sig || !sig_off || !sig_len)) return ra__fail(RA_EINVAL, "NULL signal table");
	// float form: per round only the chunk each active read is at goes up (sig_len[r] IS l_sig, rsig.cpp:213-222 ran in the caller's
	// reader), so a read that maps in its first chunk costs 16 KB of PCIe instead of its whole signal; nothing stays resident
	const bool lazy = !from_raw && !resident && !ra__map_sig_whole(ctx);
	const bool count_cells = stats && stats->reserved == 1;        // instrumentation: DP cells of the chain evaluations (costs a counter pass)
	if (ix->w != 0 || ix->n != 0) return ra__fail(RA_EUNSUP, "minimizer (w=%d) / BLEND (n=%d) seeding is not implemented on the device", ix->w, ix->n);
	if (opt->num_best_chains < 1 || opt->num_best_chains > (uint32_t)kMaxBest) return ra__fail(RA_EUNSUP, "num_best_chains %u (supported: 1..%d)", opt->num_best_chains, kMaxBest);
	if (opt->chunk_size == 0 || opt->max_num_chunk == 0) return ra__fail(RA_EINVAL, "chunk_size / max_num_chunk must be positive");
	if ((uint64_t)events_bound(opt->chunk_size, opt->window_length1, opt->window_length2) * opt->max_num_chunk >= (1ull << kQueryBits))
		return ra__fail(RA_EUNSUP, "more than %u events per read possible (chunk_size x max_num_chunk too large)", 1u << kQueryBits);
	const uint64_t cap_chains64 = (uint64_t)opt->num_best_chains * 2 * ix->n_seq;
	if (cap_chains64 > 8192) return ra__fail(RA_EUNSUP, "%u reference sequences x %u best chains exceed the per-read chain table", ix->n_seq, opt->num_best_chains);
	if (stats) memset(stats, 0, sizeof(*stats));
	if (n_reads == 0) return RA_OK;
	uint64_t cells_acc[2] = {0, 0};
	const bool cigar = (opt->flag & RA_M_DTW_OUTPUT_CIGAR) != 0;
	if (cigar && opt->dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_GLOBAL && opt->dtw_fill_method == RA_DTW_FILL_METHOD_BANDED)
		return ra__fail(RA_EUNSUP, "global border + banded fill + cigar is not implemented by the reference (rmap.cpp:223-225 asserts)");
	std::vector<std::pair<uint32_t, float>> cig_alns;
	ix->cig_start.assign(n_reads, 0); ix->cig_len.assign(n_reads, 0); ix->cig_add_i.assign(n_reads, 0); ix->cig_add_j.assign(n_reads, 0); ix->cig_total = 0;
	int rc;
	{   // hot-path options of the context follow the map options; fp_contract / test hooks stay as the context has them
		ra_mapopt_t o; if ((rc = ra_get_opt(ctx, &o))) return rc;
		o.window_length1 = opt->window_length1; o.window_length2 = opt->window_length2; o.threshold1 = opt->threshold1; o.threshold2 = opt->threshold2;
		o.peak_height = opt->peak_height; o.dtw_border_constraint = opt->dtw_border_constraint; o.dtw_fill_method = opt->dtw_fill_method;
		o.dtw_band_radius_frac = opt->dtw_band_radius_frac; o.dtw_match_bonus = opt->dtw_match_bonus; o.dtw_min_score = opt->dtw_min_score;
		if ((rc = ra_set_opt(ctx, &o))) return rc;
	}
	float *ref_dev = nullptr; uint64_t ref_n = 0;
	if ((rc = ra_ref_device(ctx, &ref_dev, &ref_n))) return rc;
	MAP_CUDA(cudaSetDevice(ra__device(ctx)));
	cudaStream_t S = ix->stream;
	StageClock clk, total; total.start(); clk.start();
	ra_map_stats_t stt; memset(&stt, 0, sizeof(stt));
	uint64_t launches = 0, lib_launches = 0;       // kernels launched here / by the K1 and DTW entry points (those count themselves)

	Params P; memset(&P, 0, sizeof(P));
	P.chunk_size = opt->chunk_size; P.max_num_chunk = opt->max_num_chunk; P.min_events = opt->min_events;
	P.max_gap_length = (int32_t)opt->max_gap_length; P.max_target_gap_length = (int32_t)opt->max_target_gap_length;
	P.chaining_band_length = (int32_t)opt->chaining_band_length; P.max_num_skips = (int32_t)opt->max_num_skips;
	P.min_num_anchors = (int32_t)opt->min_num_anchors; P.num_best_chains = (int32_t)opt->num_best_chains;
	P.min_chaining_score = opt->min_chaining_score; P.min_chain_anchor = opt->min_chain_anchor;
	P.min_bestmap_ratio = opt->min_bestmap_ratio; P.min_meanmap_ratio = opt->min_meanmap_ratio; P.dtw_min_score = opt->dtw_min_score;
	P.sample_rate = opt->sample_rate; P.bp_per_sec = opt->bp_per_sec;
	P.dtw = (opt->flag & RA_M_DTW_EVALUATE_CHAINS) ? 1u : 0u; P.no_filter = (opt->flag & RA_M_DISABLE_CHAININGSCORE_FILTERING) ? 1u : 0u;
	P.e = ix->e; P.q = ix->q; P.lq = ix->lq; P.n_seq = ix->n_seq; P.cap_chains = (uint32_t)cap_chains64; P.slot_bits = ix->slot_bits;

	// ---- signal to the device, read table ------------------------------------------------------------------------------
	if (!resident) {
		ix->last_off.assign(from_raw ? raw_off : sig_off, (from_raw ? raw_off : sig_off) + n_reads);
		ix->last_len.assign(from_raw ? raw_len : sig_len, (from_raw ? raw_len : sig_len) + n_reads);
		ix->last_raw = from_raw;
	}
	const uint64_t *in_off = ix->last_off.data();
	const uint32_t *in_len = ix->last_len.data();
	uint64_t n_samples = 0;
	for (uint32_t r = 0; r < n_reads; ++r) n_samples = std::max<uint64_t>(n_samples, in_off[r] + in_len[r]);
	std::vector<ReadState> hst(n_reads);
	uint64_t ev_cap = 0;
	const uint64_t max_read = (uint64_t)opt->chunk_size * opt->max_num_chunk;
	for (uint32_t r = 0; r < n_reads; ++r) {
		ReadState s; memset(&s, 0, sizeof(s));
		s.sig_off = in_off[r]; s.l_sig = in_len[r]; s.ev_base = ev_cap;
		const uint64_t used = std::min<uint64_t>(in_len[r], max_read);
		const uint64_t nch = (used + opt->chunk_size - 1) / opt->chunk_size;
		ev_cap += used / (opt->window_length1 / 2 + 2) + used / (opt->window_length2 / 2 + 2) + 4 * nch + 4;
		hst[r] = s;
	}
	if (!lazy && (rc = ix->sig.reserve(std::max<uint64_t>(n_samples, 1) * 4 + 64))) return rc;
	if ((rc = ix->state.reserve((size_t)n_reads * sizeof(ReadState)))) return rc;
	if ((rc = ix->chains.reserve((size_t)n_reads * P.cap_chains * sizeof(ChainRec)))) return rc;
	if ((rc = ix->paf.reserve((size_t)n_reads * sizeof(ra_paf_t)))) return rc;
	if ((rc = ix->ev.reserve(std::max<uint64_t>(ev_cap, 1) * 4))) return rc;
	if ((rc = ix->act.reserve((size_t)n_reads * 4)) || (rc = ix->coff.reserve((size_t)n_reads * 8)) || (rc = ix->clen.reserve((size_t)n_reads * 4))) return rc;
	if ((rc = ix->counters.reserve(256))) return rc;
	if ((rc = ix->evoff_round.reserve(((size_t)n_reads + 1) * 8))) return rc;
	if ((rc = ix->n_seed.reserve((size_t)n_reads * 4)) || (rc = ix->seg_n.reserve((size_t)n_reads * 4)) || (rc = ix->n_hits.reserve((size_t)n_reads * 4)) ||
	    (rc = ix->scan_tmp.reserve((size_t)n_reads * 4)) || (rc = ix->seg_base.reserve(((size_t)n_reads + 1) * 8)) ||
	    (rc = ix->n_cand.reserve((size_t)n_reads * 4)) || (rc = ix->n_nt.reserve((size_t)n_reads * 4)) || (rc = ix->cand_base.reserve(((size_t)n_reads + 1) * 8)) ||
	    (rc = ix->group_first.reserve(((size_t)n_reads + 1) * 4)) || (rc = ix->cand.reserve((size_t)n_reads * P.cap_chains * sizeof(ChainRec)))) return rc;
	if (!ix->pin) { MAP_CUDA(cudaMallocHost(&ix->pin, 4096)); ix->pin_cap = 4096; }
	MAP_CUDA(cudaMemcpyAsync(ix->state.p, hst.data(), (size_t)n_reads * sizeof(ReadState), cudaMemcpyHostToDevice, S));
	const uint32_t *l_sig_dev = nullptr;
	if (resident) {
		if (ix->last_raw) l_sig_dev = ix->lsig_keep.as<uint32_t>();
	} else if (from_raw) {
		if ((rc = ix->raw16.reserve(std::max<uint64_t>(n_samples, 1) * 2 + 64)) || (rc = ix->rd_off.reserve((size_t)n_reads * 8)) || (rc = ix->rd_len.reserve((size_t)n_reads * 4)) ||
		    (rc = ix->rd_offset.reserve((size_t)n_reads * 4)) || (rc = ix->rd_scale.reserve((size_t)n_reads * 4)) || (rc = ix->seed_cnt.reserve((size_t)n_reads * 4))) return rc;
		MAP_CUDA(cudaMemcpyAsync(ix->raw16.p, raw, n_samples * 2, cudaMemcpyHostToDevice, S));
		MAP_CUDA(cudaMemcpyAsync(ix->rd_off.p, raw_off, (size_t)n_reads * 8, cudaMemcpyHostToDevice, S));
		MAP_CUDA(cudaMemcpyAsync(ix->rd_len.p, raw_len, (size_t)n_reads * 4, cudaMemcpyHostToDevice, S));
		MAP_CUDA(cudaMemcpyAsync(ix->rd_offset.p, offset, (size_t)n_reads * 4, cudaMemcpyHostToDevice, S));
		MAP_CUDA(cudaMemcpyAsync(ix->rd_scale.p, scale, (size_t)n_reads * 4, cudaMemcpyHostToDevice, S));
		if ((rc = ix->lsig_keep.reserve((size_t)n_reads * 4))) return rc;
		uint32_t *lsd = ix->lsig_keep.as<uint32_t>();              // per-read l_sig of the decode (kept for resident re-runs)
		MAP_CUDA(ra::launch_decode_raw(ix->raw16.as<int16_t>(), ix->rd_off.as<uint64_t>(), ix->rd_len.as<uint32_t>(), ix->rd_offset.as<float>(),
		                               ix->rd_scale.as<float>(), n_reads, ix->sig.as<float>(), lsd, S));
		l_sig_dev = lsd; ++launches;
		stt.h2d_bytes += n_samples * 2 + (uint64_t)n_reads * 20;
	} else if (!lazy) { MAP_CUDA(cudaMemcpyAsync(ix->sig.p, sig, n_samples * 4, cudaMemcpyHostToDevice, S)); stt.h2d_bytes += n_samples * 4; }
	const float *hsig_dev = nullptr;                               // the caller's buffer as the device sees it (page-locked, mapped), else NULL
	const int host_gather = ra__map_host_gather(ctx);
	if (lazy && n_samples && !host_gather) {
		cudaPointerAttributes pa0, pa1;
		if (cudaPointerGetAttributes(&pa0, sig) == cudaSuccess && cudaPointerGetAttributes(&pa1, sig + n_samples - 1) == cudaSuccess &&
		    pa0.type == cudaMemoryTypeHost && pa1.type == cudaMemoryTypeHost && pa0.devicePointer &&
		    (const char *)pa1.devicePointer - (const char *)pa0.devicePointer == (const char *)(sig + n_samples - 1) - (const char *)sig)
			hsig_dev = static_cast<const float *>(pa0.devicePointer);
		(void)cudaGetLastError();
	}
	const uint32_t lazy_stride = (opt->chunk_size + 3u) & ~3u;
	std::vector<uint64_t> h_coff; std::vector<uint32_t> h_clen;
	if (!resident) stt.h2d_bytes += (uint64_t)n_reads * sizeof(ReadState);
	k_init_state<<<(n_reads + 255) / 256, 256, 0, S>>>(ix->state.as<ReadState>(), n_reads, l_sig_dev, ix->paf.as<ra_paf_t>()); ++launches;
	MAP_CUDA(cudaStreamSynchronize(S));
	stt.decode_ms = clk.lap();

	ReadState *st = ix->state.as<ReadState>();
	uint32_t *counters = ix->counters.as<uint32_t>();
	unsigned long long *totals = reinterpret_cast<unsigned long long *>(counters + 8);
	uint32_t *hc = reinterpret_cast<uint32_t *>(ix->pin);
	const uint32_t w1 = opt->window_length1, w2 = opt->window_length2;
	for (uint32_t round = 0; round < opt->max_num_chunk; ++round) {
		const uint32_t cur = round & 1u;
		MAP_CUDA(cudaMemsetAsync(counters, 0, 64, S));
		k_round_setup<<<1, 1024, 0, S>>>(st, n_reads, P, ix->act.as<uint32_t>(), ix->coff.as<uint64_t>(), ix->clen.as<uint32_t>(), counters); ++launches;
		MAP_CUDA(cudaMemcpyAsync(hc, counters, 64, cudaMemcpyDeviceToHost, S));   // [0] active reads, [1] longest chunk, [14..15] samples of the round
		MAP_CUDA(cudaStreamSynchronize(S));
		const uint32_t n_act = hc[0], max_len = hc[1];
		if (n_act == 0) break;
		stt.rounds = round + 1; stt.chunks += n_act;
		stt.select_ms += clk.lap();
		if (lazy) {
			if ((rc = ix->sig.reserve((uint64_t)n_act * lazy_stride * 4 + 64))) return rc;
			if (hsig_dev) {
				MAP_CUDA(cudaEventRecord(ix->ev_t[0], ix->stream_hi));
				k_gather_chunks<<<std::min<uint32_t>((n_act + 7) / 8, 148), 256, 0, ix->stream_hi>>>(hsig_dev, ix->coff.as<uint64_t>(), ix->clen.as<uint32_t>(), n_act, ix->sig.as<float>(), lazy_stride); ++launches;
				MAP_CUDA(cudaEventRecord(ix->ev_t[1], ix->stream_hi));
				stt.h2d_bytes += 4ull * *reinterpret_cast<const uint64_t *>(hc + 14);
			} else {                                                // pageable caller memory: gather on the host into a page-locked staging buffer
				h_coff.resize(n_act); h_clen.resize(n_act);
				MAP_CUDA(cudaMemcpyAsync(h_coff.data(), ix->coff.p, (size_t)n_act * 8, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(h_clen.data(), ix->clen.p, (size_t)n_act * 4, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaStreamSynchronize(S));
				const size_t need = (size_t)n_act * lazy_stride * 4;
				if (ix->pin_sig_cap < need) {
					if (ix->pin_sig) cudaFreeHost(ix->pin_sig);
					ix->pin_sig = nullptr; ix->pin_sig_cap = 0;
					MAP_CUDA(cudaMallocHost(&ix->pin_sig, need)); ix->pin_sig_cap = need;
				}
				float *stg = static_cast<float *>(ix->pin_sig);
				uint64_t moved = 0;
				const int nthr = std::max(1, std::min(host_gather, (int)(n_act / 256)));
				auto gather = [&](uint32_t i0, uint32_t i1) { for (uint32_t i = i0; i < i1; ++i) memcpy(stg + (size_t)i * lazy_stride, sig + h_coff[i], (size_t)h_clen[i] * 4); };
				if (nthr > 1) {
					std::vector<std::thread> th;
					for (int t = 1; t < nthr; ++t) th.emplace_back(gather, (uint32_t)((uint64_t)n_act * t / nthr), (uint32_t)((uint64_t)n_act * (t + 1) / nthr));
					gather(0, (uint32_t)((uint64_t)n_act / nthr));
					for (auto &t : th) t.join();
				} else gather(0, n_act);
				for (uint32_t i = 0; i < n_act; ++i) moved += h_clen[i];
				MAP_CUDA(cudaMemcpyAsync(ix->sig.p, stg, need, cudaMemcpyHostToDevice, S));
				k_dense_coff<<<(n_act + 255) / 256, 256, 0, S>>>(ix->coff.as<uint64_t>(), n_act, lazy_stride); ++launches;
				stt.h2d_bytes += 4 * moved;
				stt.d2h_bytes += (uint64_t)n_act * 12;
			}
			MAP_CUDA(cudaStreamSynchronize(hsig_dev ? ix->stream_hi : S));   // K1 runs on the context's streams: the round arena must be complete (S is idle here: the round's counters were just read)
			stt.decode_ms += clk.lap();
			if (hsig_dev && getenv("RA_MAP_TRACE")) { float gms = 0; cudaEventElapsedTime(&gms, ix->ev_t[0], ix->ev_t[1]); fprintf(stderr, "[map] round %u: gather of %u chunks %.3f ms on the device\n", round, n_act, gms); }
		}
		// ---- K1 over the round's chunks, appended to the reads' event arrays ------------------------------------------
		const uint64_t round_cap = (uint64_t)n_act * events_bound(max_len, w1, w2);
		if ((rc = ix->ev_round.reserve(round_cap * 4)) || (rc = ix->seeds_h.reserve(round_cap * 4)) || (rc = ix->seeds_p.reserve(round_cap * 4)) ||
		    (rc = ix->seed_cnt.reserve(round_cap * 4)) || (rc = ix->seed_off.reserve(round_cap * 8))) return rc;
		uint64_t tot_ev = 0;
		const uint64_t l0 = ra_launch_count(ctx);
		if ((rc = ra_detect_events_batch_dev(ctx, ix->sig.as<float>(), ix->coff.as<uint64_t>(), ix->clen.as<uint32_t>(), n_act, max_len,
		                                     ix->ev_round.as<float>(), round_cap, ix->evoff_round.as<uint64_t>(), &tot_ev))) return rc;
		stt.events += tot_ev;
		k_append_events<<<(n_act * 32 + 127) / 128, 128, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, ix->ev_round.as<float>(), ix->evoff_round.as<uint64_t>(),
		                                                      ix->ev.as<float>()); ++launches;
		MAP_CUDA(cudaStreamSynchronize(S));
		stt.events_ms += clk.lap();
		// ---- seeds, index lookups, anchor counts ------------------------------------------------------------------------
		uint32_t *skip = ix->scan_tmp.as<uint32_t>();
		uint32_t *seg_pad = ix->n_hits.as<uint32_t>();
		const size_t sym_bytes = (events_bound(max_len, w1, w2) + 63u) & ~63u;
		if (sym_bytes > 48 * 1024) return ra__fail(RA_EUNSUP, "chunk_size %u needs %zu bytes of symbols per warp", opt->chunk_size, sym_bytes);
		k_seed<<<n_act, 32, sym_bytes, S>>>(st, ix->act.as<uint32_t>(), n_act, ix->ev_round.as<float>(), ix->evoff_round.as<uint64_t>(), P, ix->slots.as<Slot>(),
		                                    ix->chains.as<ChainRec>(), ix->seeds_h.as<uint32_t>(), ix->seeds_p.as<uint32_t>(), ix->seed_cnt.as<uint32_t>(),
		                                    ix->seed_off.as<uint64_t>(), ix->n_seed.as<uint32_t>(), ix->seg_n.as<uint32_t>(), seg_pad, skip, totals); ++launches;
		MAP_CUDA(ra::launch_exclusive_scan_u32(seg_pad, ix->seg_base.as<uint64_t>(), n_act, S)); ++launches;
		uint64_t *h64 = reinterpret_cast<uint64_t *>(hc + 16);
		MAP_CUDA(cudaMemcpyAsync(h64, ix->seg_base.as<uint64_t>() + n_act, 8, cudaMemcpyDeviceToHost, S));
		MAP_CUDA(cudaMemcpyAsync(h64 + 1, totals, 24, cudaMemcpyDeviceToHost, S));
		MAP_CUDA(cudaStreamSynchronize(S));
		const uint64_t total_pad = h64[0];
		stt.seeds += h64[1]; stt.anchors += h64[2];
		const uint64_t max_n = h64[3];                                 // longest anchor segment of the round
		stt.seed_ms += clk.lap();
		// ---- anchors, sort, chaining -------------------------------------------------------------------------------------
		const uint64_t an = std::max<uint64_t>(total_pad, 1);
		if ((rc = ix->keys.reserve(an * 8)) || (rc = ix->score.reserve(an * 4)) || (rc = ix->pred.reserve(an * 4)) || (rc = ix->used.reserve(an)) || (rc = ix->nt.reserve(an * 4)) || (rc = ix->ntf.reserve(an)) || (rc = ix->cxf.reserve(an)) ||
		    (rc = ix->anc[cur].reserve(an * sizeof(ra_anchor_t)))) return rc;
		k_fill_anchors<<<n_act, 32, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, ix->evoff_round.as<uint64_t>(), P, ix->chains.as<ChainRec>(),
		                                    ix->anc[0].as<ra_anchor_t>(), ix->anc[1].as<ra_anchor_t>(), ix->anc[cur].as<ra_anchor_t>(), cur,
		                                    ix->seeds_p.as<uint32_t>(), ix->seed_cnt.as<uint32_t>(), ix->seed_off.as<uint64_t>(), ix->vals.as<uint64_t>(),
		                                    ix->n_seed.as<uint32_t>(), ix->seg_n.as<uint32_t>(), seg_pad, ix->seg_base.as<uint64_t>(), skip, ix->keys.as<uint64_t>());
		k_advance_offset<<<(n_act + 127) / 128, 128, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, ix->evoff_round.as<uint64_t>(), skip);
		if (!ix->sort_attr) {
			MAP_CUDA(cudaFuncSetAttribute(k_sort_segments, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kSortSmem * 8)));
			MAP_CUDA(cudaFuncSetAttribute(k_sort_buckets<kBucketMaxKeys, 0u>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BucketSmemT<kBucketMaxKeys>)));
			MAP_CUDA(cudaFuncSetAttribute(k_sort_buckets<kBucketMaxKeysL, kBucketMaxKeys>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BucketSmemT<kBucketMaxKeysL>)));
			MAP_CUDA(cudaFuncSetAttribute(k_sort_buckets<0u, kBucketMaxKeysL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BucketSmemT<0u>)));
			ix->sort_attr = true;
		}
		uint32_t done_n = 0;
		if (ix->n_bins && !getenv("RA_MAP_BITONIC_ONLY")) {            // bucket form for segments that fit shared memory, the bitonic network for the rest
			k_sort_buckets<kBucketMaxKeys, 0u><<<n_act, kBucketThreads, sizeof(BucketSmemT<kBucketMaxKeys>), S>>>(
			    ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), ix->seg_n.as<uint32_t>(), seg_pad, skip, n_act, ix->bin_base.as<uint32_t>(), ix->tshift, ix->n_bins, nullptr);
			++launches;
			if (max_n > kBucketMaxKeys) {                           // segments of 8193 .. 20480 anchors: same algorithm, one CTA per SM
				k_sort_buckets<kBucketMaxKeysL, kBucketMaxKeys><<<n_act, kBucketThreads, sizeof(BucketSmemT<kBucketMaxKeysL>), S>>>(
				    ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), ix->seg_n.as<uint32_t>(), seg_pad, skip, n_act, ix->bin_base.as<uint32_t>(), ix->tshift, ix->n_bins, nullptr);
				++launches;
			}
			if (max_n > kBucketMaxKeysL) {                          // longer still: the binned keys are staged in global memory
				if ((rc = ix->stage.reserve(an * 8))) return rc;
				k_sort_buckets<0u, kBucketMaxKeysL><<<n_act, kBucketThreads, sizeof(BucketSmemT<0u>), S>>>(
				    ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), ix->seg_n.as<uint32_t>(), seg_pad, skip, n_act, ix->bin_base.as<uint32_t>(), ix->tshift, ix->n_bins,
				    ix->stage.as<uint64_t>());
				++launches;
			}
			done_n = 0xffffffffu;
		}
		if (max_n > done_n)
			k_sort_segments<<<n_act, kSortThreads, kSortSmem * 8, S>>>(ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), seg_pad, skip, n_act, ix->seg_n.as<uint32_t>(), done_n);
		launches += 3;
		MAP_CUDA(cudaStreamSynchronize(S));
		stt.sort_ms += clk.lap();
		k_chain_pre<<<n_act, kPreThreads, 0, S>>>(ix->act.as<uint32_t>(), n_act, P, ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), ix->seg_n.as<uint32_t>(), skip,
		                                                     ix->score.as<float>(), ix->pred.as<uint32_t>(), ix->used.as<uint8_t>(), ix->nt.as<uint32_t>(), ix->n_nt.as<uint32_t>(),
		                                                     ix->ntf.as<uint8_t>());
		k_chain_simple<<<n_act, kPreThreads, 0, S>>>(ix->act.as<uint32_t>(), n_act, P, ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), skip, ix->score.as<float>(),
		                                             ix->pred.as<uint32_t>(), ix->nt.as<uint32_t>(), ix->n_nt.as<uint32_t>(), ix->ntf.as<uint8_t>(), ix->cxf.as<uint8_t>());
		k_chain<<<(n_act * 32 + 127) / 128, 128, 0, S>>>(ix->act.as<uint32_t>(), n_act, P, ix->keys.as<uint64_t>(), ix->seg_base.as<uint64_t>(), ix->seg_n.as<uint32_t>(), skip,
		                                                 ix->score.as<float>(), ix->pred.as<uint32_t>(), ix->used.as<uint8_t>(), ix->nt.as<uint32_t>(), ix->n_nt.as<uint32_t>(),
		                                                 ix->cxf.as<uint8_t>(), ix->anc[cur].as<ra_anchor_t>(), ix->cand.as<ChainRec>(), ix->n_cand.as<uint32_t>());
		launches += 2;
		MAP_CUDA(ra::launch_exclusive_scan_u32(ix->n_cand.as<uint32_t>(), ix->cand_base.as<uint64_t>(), n_act, S));
		launches += 2;
		MAP_CUDA(cudaMemcpyAsync(h64, ix->cand_base.as<uint64_t>() + n_act, 8, cudaMemcpyDeviceToHost, S));
		MAP_CUDA(cudaStreamSynchronize(S));
		const uint64_t n_cand_total = h64[0];
		stt.chain_ms += clk.lap();
		// ---- DTW evaluation of the candidates (the hot path) -------------------------------------------------------------
		if (P.dtw && n_cand_total) {
			if (n_cand_total >= 0xffffffffull) return ra__fail(RA_EINVAL, "too many candidate chains in one round");
			if ((rc = ix->dtw_chains.reserve(n_cand_total * sizeof(ra_chain_t))) || (rc = ix->dscore.reserve(n_cand_total * 4)) || (rc = ix->dkeep.reserve(n_cand_total))) return rc;
			k_emit_dtw_chains<<<(n_act + 1 + 127) / 128, 128, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, P, ix->cand.as<ChainRec>(), ix->n_cand.as<uint32_t>(),
			                                                          ix->cand_base.as<uint64_t>(), ix->seq_tab.as<uint64_t>(), ix->dtw_chains.as<ra_chain_t>(),
			                                                          ix->group_first.as<uint32_t>()); ++launches;
			MAP_CUDA(cudaStreamSynchronize(S));
			if ((rc = ra_align_chains_batch_dev(ctx, ix->ev.as<float>(), nullptr, ix->anc[cur].as<ra_anchor_t>(), an, ix->dtw_chains.as<ra_chain_t>(),
			                                    (uint32_t)n_cand_total, ix->group_first.as<uint32_t>(), n_act, 1, nullptr, ix->dscore.as<float>(),
			                                    ix->dkeep.as<uint8_t>(), count_cells ? &cells_acc[0] : nullptr, count_cells ? &cells_acc[1] : nullptr))) return rc;
			stt.cells_issued += cells_acc[0]; stt.cells_useful += cells_acc[1];
			stt.chains_evaluated += n_cand_total;
		}
		lib_launches += ra_launch_count(ctx) - l0;
		stt.dtw_ms += clk.lap();
		k_select<<<(n_act + 63) / 64, 64, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, P, ix->cand.as<ChainRec>(), ix->n_cand.as<uint32_t>(), ix->cand_base.as<uint64_t>(),
		                                         ix->dscore.as<float>(), ix->dkeep.as<uint8_t>(), skip, ix->chains.as<ChainRec>(), ix->anc[cur].as<ra_anchor_t>(), cur,
		                                         ix->seq_tab.as<uint64_t>(), ix->paf.as<ra_paf_t>(), counters); ++launches;
		MAP_CUDA(cudaGetLastError());
		if (cigar) {                                                // rmap.cpp:715-717 for the reads that left the loop mapped in this round
			if ((rc = ix->fin_r.reserve((size_t)n_act * 4)) || (rc = ix->fin_aoff.reserve((size_t)n_act * 8)) || (rc = ix->fin_chain.reserve((size_t)n_act * sizeof(ra_chain_t))) ||
			    (rc = ix->fin_anc.reserve(an * sizeof(ra_anchor_t)))) return rc;
			k_cigar_list<<<1, 1024, 0, S>>>(st, ix->act.as<uint32_t>(), n_act, P, ix->chains.as<ChainRec>(), ix->paf.as<ra_paf_t>(), ix->fin_r.as<uint32_t>(),
			                                ix->fin_aoff.as<uint64_t>(), counters); ++launches;
			MAP_CUDA(cudaMemcpyAsync(hc + 4, counters + 4, 8, cudaMemcpyDeviceToHost, S));
			MAP_CUDA(cudaStreamSynchronize(S));
			const uint32_t n_fin = hc[4];                           // (hc[5]: anchors of those chains, sized fin_anc already)
			if (n_fin) {
				k_cigar_pack<<<(n_fin * 32 + 127) / 128, 128, 0, S>>>(st, P, ix->chains.as<ChainRec>(), ix->anc[cur].as<ra_anchor_t>(), ix->seq_tab.as<uint64_t>(),
				                                                      ix->fin_r.as<uint32_t>(), ix->fin_aoff.as<uint64_t>(), n_fin, ix->fin_chain.as<ra_chain_t>(),
				                                                      ix->fin_anc.as<ra_anchor_t>()); ++launches;
				// align_chain(cigar = true) for the n_fin winning chains, all on the device (map_cigar.cuh)
				const int sparse = opt->dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_SPARSE;
				ra_index::CigDev &G = ix->cg;
				if ((rc = G.parts.reserve((size_t)n_fin * 4)) || (rc = G.first.reserve(((size_t)n_fin + 1) * 8)) || (rc = G.alns.reserve((size_t)n_fin * 4)) ||
				    (rc = G.coff.reserve((size_t)n_fin * 8)) || (rc = G.clen.reserve((size_t)n_fin * 4)) || (rc = G.add.reserve((size_t)n_fin * 16)) || (rc = G.cnt.reserve(64))) return rc;
				const ra_chain_t *fch = ix->fin_chain.as<ra_chain_t>(); const ra_anchor_t *fan = ix->fin_anc.as<ra_anchor_t>();
				k_cig_count<<<(n_fin + 255) / 256, 256, 0, S>>>(fch, n_fin, sparse, G.parts.as<uint32_t>()); ++launches;
				MAP_CUDA(ra::launch_exclusive_scan_u32(G.parts.as<uint32_t>(), G.first.as<uint64_t>(), n_fin, S)); ++launches;
				MAP_CUDA(cudaMemcpyAsync(h64, G.first.as<uint64_t>() + n_fin, 8, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaStreamSynchronize(S));
				const uint64_t n_items64 = h64[0];
				if (n_items64 >= 0xffffffffull) return ra__fail(RA_EINVAL, "too many alignment parts in one round");
				const uint32_t n_items = (uint32_t)n_items64;
				if ((rc = G.poff.reserve(((size_t)n_items + 1) * 8)) || (rc = G.items.reserve(std::max<size_t>(n_items, 1) * sizeof(ra_dtw_item_t))) ||
				    (rc = G.cost.reserve(std::max<size_t>(n_items, 1) * 4))) return rc;
				uint64_t total = 0;
				if (n_items) {
					if ((rc = G.slot_sz.reserve((size_t)n_items * 4)) || (rc = G.dir_sz.reserve((size_t)n_items * 4)) || (rc = G.slot_off.reserve(((size_t)n_items + 1) * 8)) ||
					    (rc = G.dir_off.reserve(((size_t)n_items + 1) * 8)) || (rc = G.plen.reserve((size_t)n_items * 4))) return rc;
					MAP_CUDA(cudaMemsetAsync(G.cnt.p, 0, 64, S));
					k_cig_items<<<(n_fin * 32 + 127) / 128, 128, 0, S>>>(fch, fan, n_fin, sparse, G.first.as<uint64_t>(), G.items.as<ra_dtw_item_t>(), G.slot_sz.as<uint32_t>(),
					                                                  G.dir_sz.as<uint32_t>(), G.cnt.as<uint32_t>() + 8); ++launches;
					MAP_CUDA(ra::launch_exclusive_scan_u32(G.slot_sz.as<uint32_t>(), G.slot_off.as<uint64_t>(), n_items, S));
					MAP_CUDA(ra::launch_exclusive_scan_u32(G.dir_sz.as<uint32_t>(), G.dir_off.as<uint64_t>(), n_items, S)); launches += 2;
					MAP_CUDA(cudaMemcpyAsync(h64, G.slot_off.as<uint64_t>() + n_items, 8, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaMemcpyAsync(h64 + 1, G.dir_off.as<uint64_t>() + n_items, 8, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaMemcpyAsync(h64 + 2, G.cnt.as<uint32_t>() + 8, 4, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaStreamSynchronize(S));
					const uint64_t slots = h64[0], dwords = h64[1]; const uint32_t maxw = (uint32_t)h64[2];
					if ((rc = G.dirs.reserve(std::max<uint64_t>(dwords, 1) * 4)) || (rc = G.si.reserve(slots * 4)) || (rc = G.sj.reserve(slots * 4)) || (rc = G.sd.reserve(slots * 4)) ||
					    (rc = G.oi.reserve(slots * 4)) || (rc = G.oj.reserve(slots * 4)) || (rc = G.od.reserve(slots * 4))) return rc;
					ra::DtwLaunch L; memset(&L, 0, sizeof(L));
					L.a_arena = ix->ev.as<float>(); L.b_arena = ref_dev; L.items = G.items.as<ra_dtw_item_t>(); L.n_items = n_items;
					L.cost = G.cost.as<float>(); L.work_counter = G.cnt.as<uint32_t>();
					if (!ix->n_sm) MAP_CUDA(cudaDeviceGetAttribute(&ix->n_sm, cudaDevAttrMultiProcessorCount, ra__device(ctx)));
					if (maxw > ra::dtw_generic_smem_ring_cap()) {
						uint32_t cap = 1; while (cap < maxw) cap <<= 1;
						if ((rc = G.ring.reserve((size_t)ix->n_sm * 16 * 3 * cap * sizeof(float)))) return rc;
						L.ring_scratch = G.ring.as<float>(); L.ring_cap = cap;
					}
					L.dirs = G.dirs.as<uint32_t>(); L.dir_off = G.dir_off.as<uint64_t>();
					L.path_i = G.si.as<uint32_t>(); L.path_j = G.sj.as<uint32_t>(); L.path_diff = G.sd.as<float>();
					L.slot_off = G.slot_off.as<uint64_t>(); L.path_len = G.plen.as<uint32_t>();
					int nl = 0;
					MAP_CUDA(ra::launch_dtw_generic(L, true, ix->n_sm, S, &nl));
					MAP_CUDA(ra::launch_exclusive_scan_u32(G.plen.as<uint32_t>(), G.poff.as<uint64_t>(), n_items, S));
					MAP_CUDA(ra::launch_path_compact(G.si.as<uint32_t>(), G.sj.as<uint32_t>(), G.sd.as<float>(), G.slot_off.as<uint64_t>(), G.items.as<ra_dtw_item_t>(),
					                                 G.plen.as<uint32_t>(), G.poff.as<uint64_t>(), n_items, G.oi.as<uint32_t>(), G.oj.as<uint32_t>(), G.od.as<float>(), S));
					launches += nl + 2;
				} else MAP_CUDA(cudaMemsetAsync(G.poff.p, 0, 8, S));
				ra_mapopt_t co; if ((rc = ra_get_opt(ctx, &co))) return rc;
				k_cig_finish<<<(n_fin * 32 + 127) / 128, 128, 0, S>>>(fch, fan, n_fin, sparse, G.first.as<uint64_t>(), G.items.as<ra_dtw_item_t>(), G.cost.as<float>(),
				                                                   G.poff.as<uint64_t>(), G.oi.as<uint32_t>(), G.oj.as<uint32_t>(), opt->dtw_match_bonus, (int)co.fp_contract,
				                                                   G.alns.as<float>(), G.coff.as<uint64_t>(), G.clen.as<uint32_t>(), G.add.as<uint64_t>()); ++launches;
				std::vector<uint32_t> fr(n_fin), clen(n_fin); std::vector<float> al(n_fin); std::vector<uint64_t> coff(n_fin), add(2 * (size_t)n_fin);
				MAP_CUDA(cudaMemcpyAsync(fr.data(), ix->fin_r.p, (size_t)n_fin * 4, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(al.data(), G.alns.p, (size_t)n_fin * 4, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(coff.data(), G.coff.p, (size_t)n_fin * 8, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(clen.data(), G.clen.p, (size_t)n_fin * 4, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(add.data(), G.add.p, (size_t)n_fin * 16, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaMemcpyAsync(h64, G.poff.as<uint64_t>() + n_items, 8, cudaMemcpyDeviceToHost, S));
				MAP_CUDA(cudaStreamSynchronize(S));
				total = h64[0];
				if ((rc = ix->cig_i.reserve((ix->cig_total + total) * 4, ix->cig_total * 4)) || (rc = ix->cig_j.reserve((ix->cig_total + total) * 4, ix->cig_total * 4)) ||
				    (rc = ix->cig_d.reserve((ix->cig_total + total) * 4, ix->cig_total * 4))) return rc;
				if (total) {                                            // the finished paths: 12 bytes per element into page-locked host arrays
					MAP_CUDA(cudaMemcpyAsync(static_cast<uint32_t *>(ix->cig_i.p) + ix->cig_total, G.oi.p, total * 4, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaMemcpyAsync(static_cast<uint32_t *>(ix->cig_j.p) + ix->cig_total, G.oj.p, total * 4, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaMemcpyAsync(static_cast<float *>(ix->cig_d.p) + ix->cig_total, G.od.p, total * 4, cudaMemcpyDeviceToHost, S));
					MAP_CUDA(cudaStreamSynchronize(S));
				}
				stt.d2h_bytes += total * 12 + (uint64_t)n_fin * 36;
				for (uint32_t k = 0; k < n_fin; ++k) {
					const uint32_t r = fr[k];
					ix->cig_start[r] = ix->cig_total + coff[k]; ix->cig_len[r] = clen[k]; ix->cig_add_i[r] = add[2 * (size_t)k]; ix->cig_add_j[r] = add[2 * (size_t)k + 1];
					cig_alns.push_back({r, al[k]});
				}
				ix->cig_total += total;
			}
			stt.select_ms += clk.lap();
		}
	}
	MAP_CUDA(cudaMemcpyAsync(out, ix->paf.p, (size_t)n_reads * sizeof(ra_paf_t), cudaMemcpyDeviceToHost, S));
	MAP_CUDA(cudaStreamSynchronize(S));
	stt.d2h_bytes += (uint64_t)n_reads * sizeof(ra_paf_t) + 64ull * stt.rounds;
	if (lazy) { ix->last_off.clear(); ix->last_len.clear(); }       // nothing resident to map again
	for (auto &x : cig_alns) out[x.first].alns = x.second;          // alns:f: prints chains[0].alignment_score after align_chain(cigar=true) has rewritten it (:306, :741)
	stt.select_ms += clk.lap();
	stt.total_ms = total.lap();
	stt.launches = launches + lib_launches;
	ra__add_launches(ctx, launches);
	if (stats) *stats = stt;
	return RA_OK;
}

// The alignments of the last ra_map_batch run with RA_M_DTW_OUTPUT_CIGAR, CSR by read (see the header).
extern "C" int ra_map_last_cigar(ra_index *ix, uint32_t n_reads, uint64_t *path_off, uint64_t *path_i, uint64_t *path_j, float *path_diff, uint64_t cap)
{
	if (!ix || !path_off) return ra__fail(RA_EINVAL, "NULL argument");
	if (ix->cig_len.size() != n_reads) return ra__fail(RA_EINVAL, "the last ra_map_batch mapped %zu reads, not %u", ix->cig_len.size(), n_reads);
	uint64_t w = 0;
	for (uint32_t r = 0; r < n_reads; ++r) { path_off[r] = w; w += ix->cig_len[r]; }
	path_off[n_reads] = w;
	if (!path_i) return RA_OK;                                      // sizing call
	if (w > cap) return ra__fail(RA_EINVAL, "path capacity %llu < %llu", (unsigned long long)cap, (unsigned long long)w);
	const uint32_t *ci = static_cast<const uint32_t *>(ix->cig_i.p), *cj = static_cast<const uint32_t *>(ix->cig_j.p);
	const float *cd = static_cast<const float *>(ix->cig_d.p);
	for (uint32_t r = 0; r < n_reads; ++r) {
		const uint64_t len = ix->cig_len[r];
		if (!len) continue;
		const uint64_t a = ix->cig_start[r];
		uint64_t *pi = path_i + path_off[r], *pj = path_j + path_off[r];
		for (uint64_t t = 0; t < len; ++t) { pi[t] = ci[a + t]; pj[t] = cj[a + t]; }
		pi[len - 1] += ix->cig_add_i[r]; pj[len - 1] += ix->cig_add_j[r];      // global border: rmap.cpp:230-233
		if (path_diff) memcpy(path_diff + path_off[r], cd + a, len * 4);
	}
	return RA_OK;
}
