// api.cu -- the C ABI of include/rawalign_b200.h: context, device arenas, host<->device staging,
// and the host-side replay of align_chain's control flow (src/rmap.cpp:181-313, :515-524).
#include "ra_common.cuh"
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <new>
#include <vector>

namespace {

thread_local char g_err[512] = "";

// RA_TRACE=1: host wall-clock marks of the sliced entry points on stderr (debugging aid)
struct Trace {
	bool on; double t0;
	static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
	Trace() : on(getenv("RA_TRACE") != nullptr), t0(now()) {}
	void mark(const char *what, int k = -1) const { if (on) fprintf(stderr, "[ra trace] %8.3f ms  %s %d\n", now() - t0, what, k); }
};

int fail(int code, const char *fmt, ...)
{
	va_list ap; va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
	return code;
}

#define RA_CUDA(call)                                                                                   \
	do {                                                                                                \
		cudaError_t e__ = (call);                                                                       \
		if (e__ != cudaSuccess) return fail(RA_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
	} while (0)

constexpr int kMaxSlices = 8;
constexpr int kCopySliceStreams = 4;  // streams the sliced host<->device entry points spread their slices over
constexpr int kSliceStreams = 8;      // chunk groups of the resident K1 (hosts should export CUDA_DEVICE_MAX_CONNECTIONS=32: more streams than hardware queues means false dependencies)
// host<->device copies of the e2e entry points are cut into slices that overlap the kernels

// grow-only device buffer
struct DevBuf {
	void *p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap) return RA_OK;
		if (p) { cudaFree(p); p = nullptr; cap = 0; }
		size_t want = bytes + bytes / 4 + 256;
		cudaError_t e = cudaMalloc(&p, want);
		if (e != cudaSuccess) { p = nullptr; cap = 0; (void)cudaGetLastError(); return fail(RA_ENOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e)); }
		cap = want;
		return RA_OK;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// grow-only pinned host buffer (staging for the small tables that cross the bus inside the sliced entry points)
struct PinBuf {
	void *p = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap) return RA_OK;
		if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
		size_t want = bytes + bytes / 4 + 256;
		cudaError_t e = cudaMallocHost(&p, want);
		if (e != cudaSuccess) { p = nullptr; cap = 0; (void)cudaGetLastError(); return fail(RA_ENOMEM, "cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e)); }
		cap = want;
		return RA_OK;
	}
	void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
	template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

} // namespace

struct ra_ctx {
	int device = 0;
	int n_sm = 0;
	cudaStream_t stream = nullptr;       // compute
	cudaStream_t h2d = nullptr, d2h = nullptr, aux = nullptr;   // copy engines and the decode kernel of the sliced e2e entry points
	cudaStream_t slice_hi[kSliceStreams] = {};                    // high-priority twins for the latency-bound K1 kernels (RA_K1_PRIO)
	cudaEvent_t split_ev[kSliceStreams][4] = {};
	cudaEvent_t grp_scan[kSliceStreams] = {}, grp_done[kSliceStreams] = {};   // per chunk group of the resident K1
	int k1_groups = 8; bool k1_prio = true;                       // A/B on one box (profiles/r01_ab_k1_groups_prio.txt); RA_K1_GROUPS / RA_K1_PRIO override
	cudaStream_t slice_stream[kSliceStreams] = {};                // K1 of different slices runs concurrently (each slice is latency-bound)
	cudaEvent_t slice_in[kMaxSlices] = {}, slice_done[kMaxSlices] = {}, slice_dec[kMaxSlices] = {}, slice_scan[kMaxSlices] = {};
	bool slices_concurrent = false;
	ra_mapopt_t opt;
	uint64_t launches = 0;
	cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // K1 start / after scan / end, DTW start / end
	bool timed_k1 = false, timed_dtw = false;
	// reference mirror
	DevBuf ref; uint64_t ref_n = 0;
	// events
	DevBuf sig, chunk_off, chunk_len, raw, nev, ev_off, ev_out, ev_pos;
	DevBuf raw16, rd_off, rd_len, rd_offset, rd_scale, rd_lsig;   // raw-signal decode (decode.cu)
	PinBuf pin_off, pin_tab, pin_lsig, pin_gf;
	uint64_t last_events = 0;
	// dtw
	DevBuf a_arena, b_arena, items, cost, counter, ring, order;
	// chains
	DevBuf anchors, chains, group_first, min_score, score, keep, cells, chain_ring;
	DevBuf gl_chain_item, gl_item_cells;                           // global-border plan (chains_global.cu)
	DevBuf sp_nparts, sp_ntiles, sp_pbase, sp_tbase, sp_tchain, sp_cost, sp_pcells, sp_best, sp_rest;   // sparse-border plan (chains_sparse.cu)
	uint64_t resident_events = 0;      // events left on the device by the last ra_detect_events_batch
	bool force_unfused_events = false; // test hook: run the five-kernel K1 v2 (also the path for windows > kFusedMaxWindow)
	bool force_fused_events = false;   // test hook: run the single-pass K1 v3
	bool force_generic_dtw = false;    // test hook (reserved bit 2): every item through the generic shared-memory DTW kernel
	bool legacy_sparse_tiles = false;  // test hook (reserved bit 3): sparse parts through the CTA-wide tiles of k_sp_parts instead of k_sp_parts2
	bool want_ev_pos = true;           // keep the event start samples of the last K1 call (ra_last_event_positions)
	int map_host_gather = 0;           // ra_set_tuning("map_host_gather"): float-form ra_map_batch gathers the round's chunks on the host (N threads) and uploads them by DMA
	bool map_sig_whole = false;        // ra_set_tuning("map_sig_whole"): float-form ra_map_batch uploads the whole batch up front
	uint64_t chain_slice_min_anchors = 2u << 20;   // ra_set_tuning("chain_slice_min_anchors"): anchors per upload slice of ra_align_chains_batch
	// traceback
	DevBuf dirs, dir_off, slot_off, slot_i, slot_j, slot_d, path_len, path_off, out_i, out_j, out_d;
};

namespace {

// Scratch for every chunk of the batch at once (groups of chunks then run on their own streams): worth it while it stays small.
constexpr uint64_t kConcurrentSliceChunks = 65536;
constexpr uint32_t kEventSubBatch = 32768;   // chunks per launch group: bounds the scratch arena (~2.4 GB at 4000 samples)

// Enqueues K1 for chunks [0, n_chunks) of the given tables on the compute stream; no host synchronisation.
// chained_first != 0 continues the CSR offsets from the value already in ev_off_dev[0].
int events_launch(ra_ctx *ctx, const float *sig_dev, const uint64_t *chunk_off_dev, const uint32_t *chunk_len_dev,
                  uint32_t n_chunks, uint32_t max_chunk_len, float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev,
                  int chained_first, bool record_timing)
{
	const ra_mapopt_t &o = ctx->opt;
	if (!ra::events_windows_supported(o.window_length1, o.window_length2))
		return fail(RA_EUNSUP, "event detector windows (%u,%u) are not supported", o.window_length1, o.window_length2);
	if (n_chunks == 0) {
		if (!chained_first) RA_CUDA(cudaMemsetAsync(ev_off_dev, 0, 8, ctx->stream));
		return RA_OK;
	}
	ra::EvParams p = { o.window_length1, o.window_length2, o.threshold1, o.threshold2, o.peak_height };
	// K1 variants: v4 pipeline (default), v3 fused single pass (reserved bit 1), v2 five kernels (reserved bit 0 or windows > 14)
	const bool pipe = ra::events_pipe_supported(o.window_length1, o.window_length2) && !ctx->force_unfused_events && !ctx->force_fused_events;
	const bool fused = !pipe && ra::events_fused_supported(o.window_length1, o.window_length2) && !ctx->force_unfused_events;
	ra::EvScratch lay; ra::events_scratch_layout(max_chunk_len, p, &lay);
	size_t fpc = lay.floats_per_chunk;
	if (fused) { uint32_t stride; ra::events_fused_layout(max_chunk_len, p, &stride, &fpc); }
	if (pipe) { uint32_t stride, pstride; ra::events_pipe_layout(max_chunk_len, p, &stride, &pstride, &fpc); }
	const uint32_t sub = std::min(n_chunks, kEventSubBatch);
	int rc;
	if (ctx->raw.cap < (size_t)sub * fpc * sizeof(float) || ctx->nev.cap < (size_t)sub * sizeof(uint32_t) ||
	    ctx->ev_pos.cap < (size_t)std::max<uint64_t>(ev_cap, 1) * sizeof(uint32_t))
		RA_CUDA(cudaStreamSynchronize(ctx->stream));             // a grow-only buffer is about to be reallocated: drain its users first
	if ((rc = ctx->raw.reserve((size_t)sub * fpc * sizeof(float)))) return rc;
	if ((rc = ctx->nev.reserve((size_t)sub * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->ev_pos.reserve((size_t)std::max<uint64_t>(ev_cap, 1) * sizeof(uint32_t)))) return rc;
	if (record_timing) RA_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
	for (uint32_t s0 = 0; s0 < n_chunks; s0 += sub) {
		const uint32_t n = std::min(sub, n_chunks - s0);
		const bool last = (s0 + n == n_chunks);
		const int chained = (s0 != 0 || chained_first) ? 1 : 0;
		cudaEvent_t after = (last && record_timing) ? ctx->ev[1] : nullptr;
		if (pipe) {
			RA_CUDA(ra::launch_events_pipe(sig_dev, chunk_off_dev + s0, chunk_len_dev + s0, n, max_chunk_len, p, (int)o.fp_contract,
			                               ctx->raw.as<float>(), ctx->nev.as<uint32_t>(), ev_off_dev + s0, chained, ev_out_dev, ev_cap,
			                               ctx->want_ev_pos ? ctx->ev_pos.as<uint32_t>() : nullptr, ctx->stream, after));
			ctx->launches += ra::kPipeEventKernels;
		} else if (fused) {
			RA_CUDA(ra::launch_events_fused(sig_dev, chunk_off_dev + s0, chunk_len_dev + s0, n, max_chunk_len, p, (int)o.fp_contract,
			                                ctx->raw.as<float>(), ctx->nev.as<uint32_t>(), ev_off_dev + s0, chained, ev_out_dev, ev_cap,
			                                ctx->want_ev_pos ? ctx->ev_pos.as<uint32_t>() : nullptr, ctx->n_sm, ctx->stream, after));
			ctx->launches += ra::kFusedEventKernels;
		} else {
			RA_CUDA(ra::launch_events(sig_dev, chunk_off_dev + s0, chunk_len_dev + s0, n, max_chunk_len, p, (int)o.fp_contract,
			                          ctx->raw.as<float>(), ctx->nev.as<uint32_t>(), ev_off_dev + s0, chained, ev_out_dev, ev_cap,
			                          ctx->ev_pos.as<uint32_t>(), ctx->stream, after));
			ctx->launches += ra::kEventKernels;
		}
	}
	if (record_timing) { RA_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream)); ctx->timed_k1 = true; }
	return RA_OK;
}

// Device-resident K1.  The stages of the v4 pipeline are bound differently (t-stat: issue; prefix sums and the detector
// state machine: dependent-instruction latency), so large batches are cut into k1_groups (8) groups of chunks that run on
// their own streams and fill each other's stalls; only the CSR scans are chained (group k waits for the scan of group k-1).
int events_dev(ra_ctx *ctx, const float *sig_dev, const uint64_t *chunk_off_dev, const uint32_t *chunk_len_dev,
               uint32_t n_chunks, uint32_t max_chunk_len, float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev,
               uint64_t *total_events)
{
	const ra_mapopt_t &o = ctx->opt;
	const bool pipe = ra::events_pipe_supported(o.window_length1, o.window_length2) && !ctx->force_unfused_events && !ctx->force_fused_events;
	int rc;
	if (pipe && n_chunks >= 2048u && n_chunks <= kConcurrentSliceChunks) {
		ra::EvParams p = { o.window_length1, o.window_length2, o.threshold1, o.threshold2, o.peak_height };
		uint32_t stride, pstride; size_t fpc;
		ra::events_pipe_layout(max_chunk_len, p, &stride, &pstride, &fpc);
		if ((rc = ctx->raw.reserve((size_t)n_chunks * fpc * sizeof(float)))) return rc;
		if ((rc = ctx->nev.reserve((size_t)n_chunks * sizeof(uint32_t)))) return rc;
		if ((rc = ctx->ev_pos.reserve((size_t)std::max<uint64_t>(ev_cap, 1) * sizeof(uint32_t)))) return rc;
		RA_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
		const int G = ctx->k1_groups;
		static cudaEvent_t tr_ev[kSliceStreams][5], tr_t0; static bool tr_init = false;
		const bool k1_trace = getenv("RA_K1_TRACE") != nullptr;
		if (k1_trace && !tr_init) { for (auto &g : tr_ev) for (auto &e : g) cudaEventCreate(&e); cudaEventCreate(&tr_t0); tr_init = true; }
		if (k1_trace) cudaEventRecord(tr_t0, ctx->stream);
		for (int k = 0; k < G; ++k) {
			ra::g_pipe_trace = k1_trace ? tr_ev[k] : nullptr;
			const uint32_t c0 = (uint32_t)((uint64_t)n_chunks * k / G), c1 = (uint32_t)((uint64_t)n_chunks * (k + 1) / G);
			cudaStream_t s = ctx->slice_stream[k];
			RA_CUDA(cudaStreamWaitEvent(s, ctx->ev[0], 0));
			ra::PipeSplit sp = { ctx->slice_hi[k], { ctx->split_ev[k][0], ctx->split_ev[k][1], ctx->split_ev[k][2], ctx->split_ev[k][3] } };
			RA_CUDA(ra::launch_events_pipe(sig_dev, chunk_off_dev + c0, chunk_len_dev + c0, c1 - c0, max_chunk_len, p, (int)o.fp_contract,
			                               ctx->raw.as<float>() + (size_t)c0 * fpc, ctx->nev.as<uint32_t>() + c0, ev_off_dev + c0, k != 0, ev_out_dev, ev_cap,
			                               ctx->want_ev_pos ? ctx->ev_pos.as<uint32_t>() : nullptr, s, nullptr,
			                               k ? ctx->grp_scan[k - 1] : nullptr, ctx->grp_scan[k], ctx->k1_prio ? &sp : nullptr));
			ctx->launches += ra::kPipeEventKernels;
			RA_CUDA(cudaEventRecord(ctx->grp_done[k], s));
		}
		ra::g_pipe_trace = nullptr;
		for (int k = 0; k < G; ++k) RA_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->grp_done[k], 0));
		if (k1_trace) {                                             // timeline of the groups' stages (ms after the start), debug aid
			cudaStreamSynchronize(ctx->stream);
			for (int k = 0; k < G; ++k) {
				float t[5] = {0, 0, 0, 0, 0};
				for (int j = 0; j < 5; ++j) cudaEventElapsedTime(&t[j], tr_t0, tr_ev[k][j]);
				fprintf(stderr, "[k1 trace] group %d: prefix %.3f tstat %.3f fsm %.3f scan %.3f finish %.3f\n", k, t[0], t[1], t[2], t[3], t[4]);
			}
		}
		RA_CUDA(cudaEventRecord(ctx->ev[1], ctx->stream));
		RA_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
		ctx->timed_k1 = true;
	} else if ((rc = events_launch(ctx, sig_dev, chunk_off_dev, chunk_len_dev, n_chunks, max_chunk_len, ev_out_dev, ev_cap, ev_off_dev, 0, true))) return rc;
	uint64_t tot = 0;
	RA_CUDA(cudaMemcpyAsync(&tot, ev_off_dev + n_chunks, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	if (total_events) *total_events = tot;
	ctx->last_events = tot;
	if (tot > ev_cap) return fail(RA_EINVAL, "ev_out capacity %llu < %llu events", (unsigned long long)ev_cap, (unsigned long long)tot);
	return RA_OK;
}

// Sliced K1: slice k = chunks [c0, c0+n) runs as soon as its inputs are on the device (slice_in[k]); the CSR offsets of
// every finished slice go to pinned staging, and slices_collect streams the events of finished slices back to the host
// on the d2h stream while later slices still compute.
int slice_launch_concurrent(ra_ctx *ctx, const float *sig_dev, int k, uint32_t c0, uint32_t n, uint32_t max_len, uint64_t dev_cap);

int slice_launch(ra_ctx *ctx, const float *sig_dev, int k, uint32_t c0, uint32_t n, uint32_t max_len, uint64_t dev_cap)
{
	int rc;
	if (ctx->slices_concurrent) return slice_launch_concurrent(ctx, sig_dev, k, c0, n, max_len, dev_cap);
	RA_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->slice_in[k], 0));
	if ((rc = events_launch(ctx, sig_dev, ctx->chunk_off.as<uint64_t>() + c0, ctx->chunk_len.as<uint32_t>() + c0, n, max_len,
	                        ctx->ev_out.as<float>(), dev_cap, ctx->ev_off.as<uint64_t>() + c0, k != 0, k == 0))) return rc;
	uint64_t *stage = ctx->pin_off.as<uint64_t>();
	RA_CUDA(cudaMemcpyAsync(stage + c0 + (k ? 1 : 0), ctx->ev_off.as<uint64_t>() + c0 + (k ? 1 : 0), (size_t)(n + (k ? 0 : 1)) * sizeof(uint64_t),
	                        cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaEventRecord(ctx->slice_done[k], ctx->stream));
	return RA_OK;
}

// Reserves the K1 scratch of the v4 pipeline for n_total chunks and decides whether slices can run concurrently.
int slices_prepare(ra_ctx *ctx, uint32_t n_total, uint32_t max_len, uint64_t dev_cap)
{
	const ra_mapopt_t &o = ctx->opt;
	const bool pipe = ra::events_pipe_supported(o.window_length1, o.window_length2) && !ctx->force_unfused_events && !ctx->force_fused_events;
	ctx->slices_concurrent = pipe && n_total <= kConcurrentSliceChunks;
	if (!ctx->slices_concurrent) return RA_OK;
	ra::EvParams p = { o.window_length1, o.window_length2, o.threshold1, o.threshold2, o.peak_height };
	uint32_t stride, pstride; size_t fpc;
	ra::events_pipe_layout(max_len, p, &stride, &pstride, &fpc);
	int rc;
	if ((rc = ctx->raw.reserve((size_t)std::max<uint32_t>(n_total, 1) * fpc * sizeof(float)))) return rc;
	if ((rc = ctx->nev.reserve((size_t)std::max<uint32_t>(n_total, 1) * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->ev_pos.reserve((size_t)std::max<uint64_t>(dev_cap, 1) * sizeof(uint32_t)))) return rc;
	return RA_OK;
}

// RA_TRACE: device-side timestamps of slice k (start of K1, end of finish kernel, end of the offsets copy)
cudaEvent_t g_tr_ev[kMaxSlices][3]; cudaEvent_t g_tr_t0 = nullptr; bool g_tr_init = false;
void trace_events_init()
{
	if (g_tr_init) return;
	cudaEventCreate(&g_tr_t0);
	for (int k = 0; k < kMaxSlices; ++k) for (int j = 0; j < 3; ++j) cudaEventCreate(&g_tr_ev[k][j]);
	g_tr_init = true;
}

// Slice k on its own stream: detection kernels run as soon as the slice's inputs are there; only the CSR scan waits for
// the previous slice's scan.
int slice_launch_concurrent(ra_ctx *ctx, const float *sig_dev, int k, uint32_t c0, uint32_t n, uint32_t max_len, uint64_t dev_cap)
{
	const ra_mapopt_t &o = ctx->opt;
	cudaStream_t s = ctx->slice_stream[k % kCopySliceStreams];
	RA_CUDA(cudaStreamWaitEvent(s, ctx->slice_in[k], 0));
	const bool trace = getenv("RA_TRACE") != nullptr;
	if (trace) { trace_events_init(); cudaEventRecord(g_tr_ev[k][0], s); }
	if (k == 0) RA_CUDA(cudaEventRecord(ctx->ev[0], s));
	uint64_t *ev_off_dev = ctx->ev_off.as<uint64_t>() + c0;
	if (n) {
		ra::EvParams p = { o.window_length1, o.window_length2, o.threshold1, o.threshold2, o.peak_height };
		uint32_t stride, pstride; size_t fpc;
		ra::events_pipe_layout(max_len, p, &stride, &pstride, &fpc);
		RA_CUDA(ra::launch_events_pipe(sig_dev, ctx->chunk_off.as<uint64_t>() + c0, ctx->chunk_len.as<uint32_t>() + c0, n, max_len, p, (int)o.fp_contract,
		                               ctx->raw.as<float>() + (size_t)c0 * fpc, ctx->nev.as<uint32_t>() + c0, ev_off_dev, k != 0, ctx->ev_out.as<float>(), dev_cap,
		                               ctx->want_ev_pos ? ctx->ev_pos.as<uint32_t>() : nullptr, s, k == 0 ? ctx->ev[1] : nullptr,
		                               k ? ctx->slice_scan[k - 1] : nullptr, ctx->slice_scan[k]));
		ctx->launches += ra::kPipeEventKernels;
	} else {
		// empty slice: pass the running total on (ev_off_dev[0] already holds it once the previous scan is done)
		if (k) RA_CUDA(cudaStreamWaitEvent(s, ctx->slice_scan[k - 1], 0));
		else RA_CUDA(cudaMemsetAsync(ev_off_dev, 0, 8, s));
		RA_CUDA(cudaEventRecord(ctx->slice_scan[k], s));
	}
	uint64_t *stage = ctx->pin_off.as<uint64_t>();
	if (trace) cudaEventRecord(g_tr_ev[k][1], s);
	RA_CUDA(cudaMemcpyAsync(stage + c0 + (k ? 1 : 0), ev_off_dev + (k ? 1 : 0), (size_t)(n + (k ? 0 : 1)) * sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
	if (trace) cudaEventRecord(g_tr_ev[k][2], s);
	RA_CUDA(cudaEventRecord(ctx->slice_done[k], s));
	return RA_OK;
}

int slices_collect(ra_ctx *ctx, const uint32_t *cut, int n_slices, uint64_t dev_cap, float *ev_out, uint64_t ev_cap, uint64_t *ev_off,
                   uint32_t n_chunks)
{
	if (ctx->slices_concurrent) {
		for (int k = 0; k < n_slices; ++k) RA_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->slice_done[k], 0));   // later work on the compute stream sees the events
		RA_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
		ctx->timed_k1 = true;
	} else RA_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
	const uint64_t *stage = ctx->pin_off.as<uint64_t>();
	Trace tr;
	for (int k = 0; k < n_slices; ++k) {
		RA_CUDA(cudaEventSynchronize(ctx->slice_done[k]));
		tr.mark("  slice done", k);
		const uint64_t e0 = stage[cut[k]], e1 = stage[cut[k + 1]];
		if (e1 > dev_cap || e1 > ev_cap) { cudaStreamSynchronize(ctx->stream); return fail(RA_EINVAL, "ev_out capacity %llu < %llu events", (unsigned long long)std::min(dev_cap, ev_cap), (unsigned long long)e1); }
		if (e1 > e0) RA_CUDA(cudaMemcpyAsync(ev_out + e0, ctx->ev_out.as<float>() + e0, (size_t)(e1 - e0) * sizeof(float), cudaMemcpyDeviceToHost, ctx->d2h));
	}
	memcpy(ev_off, stage, (size_t)(n_chunks + 1) * sizeof(uint64_t));
	RA_CUDA(cudaStreamSynchronize(ctx->d2h));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->last_events = ctx->resident_events = stage[n_chunks];
	return RA_OK;
}

// width (in floats) of the widest anti-diagonal of one item, as the generic kernel sees it
uint32_t item_width(const ra_dtw_item_t &it, bool tb)
{
	uint32_t n = std::max(it.a_len, it.b_len), m = std::min(it.a_len, it.b_len);
	if (tb) return it.b_len;
	if (it.band_radius < 0) return m;
	if (n == 0) return 0;
	const uint32_t extra = ((n - m) * (uint32_t)it.band_radius + n - 1u) / n;
	return (uint32_t)it.band_radius + extra + 1u;
}

// Launches the cost kernels for items already on the device.  items_host is used for sizing and for sorting the items into
// kernel classes: problems whose band fits one warp's register wavefront go to dtw_wave.cu (by rows per lane), the rest
// (very tall bands or full matrices) to the generic shared-memory kernel.
int dtw_dev(ra_ctx *ctx, const float *a_dev, const float *b_dev, const ra_dtw_item_t *items_dev,
            const ra_dtw_item_t *items_host, uint32_t n_items, float *cost_dev)
{
	if (n_items == 0) return RA_OK;
	int rc;
	static const int kNC = 9;
	static const int kClassH[kNC] = { 1, 2, 4, 8, 16, 32, 64 + 8, 64 + 16, 64 + 32 };   // dtw_wave_class codes
	// Two flat passes (class of every item, then a counting-sort scatter): item streams run to millions of tiny problems, and
	// per-class vectors cost more host time than the kernels take.  Slots: wave classes [0, kNC), generic kNC, then the
	// R' <= 2 items (one per thread) by height -- kTinyBins slots, so the lanes of a warp run equally long.
	constexpr uint32_t kTinyBins = 1024, kSlots = kNC + 1 + kTinyBins;
	std::vector<uint16_t> slot(n_items);
	std::vector<uint32_t> start(kSlots + 1, 0);
	uint32_t maxw = 0;
	for (uint32_t k = 0; k < n_items; ++k) {
		const ra_dtw_item_t &it = items_host[k];
		uint32_t sl;
		if (!ctx->force_generic_dtw && ra::dtw_tiny_supported(it.a_len, it.b_len, it.band_radius))
			sl = kNC + 1 + std::min<uint32_t>(std::max(it.a_len, it.b_len), kTinyBins - 1);
		else {
			const int H = ctx->force_generic_dtw ? 0 : ra::dtw_wave_class(it.a_len, it.b_len, it.band_radius);
			sl = kNC;
			for (int q = 0; q < kNC; ++q) if (H == kClassH[q]) sl = (uint32_t)q;
			if (sl == (uint32_t)kNC) maxw = std::max(maxw, item_width(it, false));
		}
		slot[k] = (uint16_t)sl;
		++start[sl + 1];
	}
	for (uint32_t q = 0; q < kSlots; ++q) start[q + 1] += start[q];
	uint32_t first[kNC + 2];
	for (int c = 0; c <= kNC + 1; ++c) first[c] = start[c];
	const uint32_t tiny_first = start[kNC + 1], n_tiny = n_items - tiny_first;
	std::vector<uint32_t> order(n_items);
	for (uint32_t k = 0; k < n_items; ++k) order[start[slot[k]]++] = k;     // stable: input order within a class
	if ((rc = ctx->order.reserve((size_t)n_items * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->counter.reserve(256))) return rc;
	RA_CUDA(cudaMemcpyAsync(ctx->order.p, order.data(), (size_t)n_items * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	ra::DtwLaunch L; memset(&L, 0, sizeof(L));
	L.a_arena = a_dev; L.b_arena = b_dev; L.items = items_dev; L.cost = cost_dev;
	RA_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
	for (int c = 0; c < kNC; ++c) {
		const uint32_t nc = first[c + 1] - first[c];
		if (!nc) continue;
		L.n_items = n_items;
		RA_CUDA(ra::launch_dtw_wave(L, kClassH[c], ctx->order.as<uint32_t>() + first[c], nc, ctx->counter.as<uint32_t>() + 6 + c, ctx->n_sm, ctx->stream));
		ctx->launches += 1;
	}
	if (n_tiny) {
		L.n_items = n_items;
		RA_CUDA(ra::launch_dtw_tiny(L, ctx->order.as<uint32_t>() + tiny_first, n_tiny, ctx->stream));
		ctx->launches += 1;
	}
	const uint32_t ng = first[kNC + 1] - first[kNC];
	if (ng) {
		L.order = ctx->order.as<uint32_t>() + first[kNC]; L.n_items = ng;
		L.work_counter = ctx->counter.as<uint32_t>();
		if (maxw > ra::dtw_generic_smem_ring_cap()) {
			uint32_t cap = 1; while (cap < maxw) cap <<= 1;
			const size_t warps = (size_t)ctx->n_sm * 4 * 4;
			if ((rc = ctx->ring.reserve(warps * 3 * cap * sizeof(float)))) return rc;
			L.ring_scratch = ctx->ring.as<float>(); L.ring_cap = cap;
		}
		int nl = 0;
		RA_CUDA(ra::launch_dtw_generic(L, false, ctx->n_sm, ctx->stream, &nl));
		ctx->launches += nl;
	}
	RA_CUDA(cudaEventRecord(ctx->ev[4], ctx->stream));
	ctx->timed_dtw = true;
	RA_CUDA(cudaStreamSynchronize(ctx->stream));                  // `order` lives on this stack frame until the copy has been consumed
	return RA_OK;
}

// Sparse border: plan (which chains, slots for their parts and tiles), flat part-cost kernel, replay -- once
// for min_score_mode 0, twice (first chain of every group, then the rest) for the running threshold.  No host
// synchronisation inside: capacities come from the anchor count (a chain of k anchors has k-1 parts).
int sparse_pipeline(ra_ctx *ctx, const ra::ChainLaunch &L, uint32_t n_chains, uint64_t n_anchors, bool reset_counters = true)
{
	int rc;
	if (n_chains == 0) return RA_OK;
	const uint32_t tile_parts = ctx->legacy_sparse_tiles ? ra::kSparseTileParts : ra::kSparseTileParts2;
	const uint64_t part_cap = n_anchors + 1, tile_cap = n_anchors / tile_parts + n_chains + 1;
	if (tile_cap >= 0xffffffffull) return fail(RA_EINVAL, "too many part tiles in one batch (%llu)", (unsigned long long)tile_cap);
	if ((rc = ctx->sp_nparts.reserve((size_t)n_chains * 4))) return rc;
	if ((rc = ctx->sp_pbase.reserve((size_t)n_chains * 8))) return rc;
	if ((rc = ctx->sp_best.reserve((size_t)L.n_groups * 4))) return rc;
	if ((rc = ctx->sp_tchain.reserve((size_t)tile_cap * 4))) return rc;
	if ((rc = ctx->sp_tbase.reserve((size_t)tile_cap * 4))) return rc;
	if ((rc = ctx->sp_cost.reserve((size_t)part_cap * 4))) return rc;
	if (L.cells && (rc = ctx->sp_pcells.reserve((size_t)part_cap * 4))) return rc;
	if ((rc = ctx->sp_ntiles.reserve(64))) return rc;
	if (!ctx->legacy_sparse_tiles && (rc = ctx->sp_rest.reserve((size_t)part_cap * 8))) return rc;
	if (reset_counters) {
		if (L.cells) RA_CUDA(cudaMemsetAsync(L.cells, 0, 2 * sizeof(uint64_t), ctx->stream));
		RA_CUDA(cudaMemsetAsync(L.need_cap, 0, sizeof(uint32_t), ctx->stream));
	}
	ra::SparsePlan P; memset(&P, 0, sizeof(P));
	P.n_chains = n_chains;
	P.n_parts = ctx->sp_nparts.as<uint32_t>(); P.part_base = ctx->sp_pbase.as<uint64_t>();
	P.tile_chain = ctx->sp_tchain.as<uint32_t>(); P.tile_t0 = ctx->sp_tbase.as<uint32_t>();
	P.totals = ctx->sp_ntiles.as<unsigned long long>();
	P.part_cost = ctx->sp_cost.as<float>(); P.part_cells = L.cells ? ctx->sp_pcells.as<uint32_t>() : nullptr;
	P.best = ctx->sp_best.as<float>();
	P.tile_parts = tile_parts; P.rest = ctx->legacy_sparse_tiles ? nullptr : ctx->sp_rest.as<uint64_t>();
	const int first = (L.min_score_mode == 1) ? 1 : 0, last = (L.min_score_mode == 1) ? 2 : 0;
	for (int phase = first; phase <= last; ++phase) {
		RA_CUDA(ra::launch_sparse_prep(L, P, phase, ctx->stream));
		RA_CUDA(ra::launch_sparse_eval(L, P, phase, ctx->n_sm, ctx->stream));
		ctx->launches += ra::kSparseKernelsPerPhase;
	}
	return RA_OK;
}

// Global border on device-resident chains: per phase, the chains whose budget passes become DTW items (k_glob_prep), the items
// visit the host only to be sorted into wavefront classes (dtw_dev), and k_glob_replay does the reference's serial arithmetic.
int global_pipeline(ra_ctx *ctx, const ra::ChainLaunch &L, uint32_t n_chains)
{
	int rc;
	if (n_chains == 0) return RA_OK;
	if ((rc = ctx->items.reserve((size_t)n_chains * sizeof(ra_dtw_item_t)))) return rc;
	if ((rc = ctx->cost.reserve((size_t)n_chains * sizeof(float)))) return rc;
	if ((rc = ctx->gl_chain_item.reserve((size_t)n_chains * sizeof(int32_t)))) return rc;
	if ((rc = ctx->sp_best.reserve((size_t)L.n_groups * sizeof(float)))) return rc;
	if (L.cells && (rc = ctx->gl_item_cells.reserve((size_t)n_chains * 8))) return rc;
	if (L.cells) RA_CUDA(cudaMemsetAsync(L.cells, 0, 2 * sizeof(uint64_t), ctx->stream));
	RA_CUDA(cudaMemsetAsync(L.need_cap, 0, sizeof(uint32_t), ctx->stream));
	uint32_t *n_items_dev = ctx->counter.as<uint32_t>() + 21;
	uint64_t issued = 0;
	const int first = (L.min_score_mode == 1) ? 1 : 0, last = (L.min_score_mode == 1) ? 2 : 0;
	std::vector<ra_dtw_item_t> items;
	std::vector<uint64_t> cells;
	for (int phase = first; phase <= last; ++phase) {
		RA_CUDA(ra::launch_glob_prep(L, phase, ctx->sp_best.as<float>(), ctx->items.as<ra_dtw_item_t>(), ctx->gl_chain_item.as<int32_t>(), n_items_dev, ctx->stream));
		uint32_t n_items = 0;
		RA_CUDA(cudaMemcpyAsync(&n_items, n_items_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));
		ctx->launches += 1;
		if (n_items) {
			items.resize(n_items);
			RA_CUDA(cudaMemcpyAsync(items.data(), ctx->items.p, (size_t)n_items * sizeof(ra_dtw_item_t), cudaMemcpyDeviceToHost, ctx->stream));
			RA_CUDA(cudaStreamSynchronize(ctx->stream));
			if (L.cells) {
				cells.resize(n_items);
				for (uint32_t k = 0; k < n_items; ++k) { cells[k] = ra_dtw_cells(items[k].a_len, items[k].b_len, items[k].band_radius); issued += cells[k]; }
				RA_CUDA(cudaMemcpyAsync(ctx->gl_item_cells.p, cells.data(), (size_t)n_items * 8, cudaMemcpyHostToDevice, ctx->stream));
			}
			if ((rc = dtw_dev(ctx, L.read_arena, L.ref_arena, ctx->items.as<ra_dtw_item_t>(), items.data(), n_items, ctx->cost.as<float>()))) return rc;
		}
		RA_CUDA(ra::launch_glob_replay(L, phase, ctx->sp_best.as<float>(), ctx->gl_chain_item.as<int32_t>(), ctx->cost.as<float>(),
		                               L.cells ? ctx->gl_item_cells.as<unsigned long long>() : nullptr, ctx->stream));
		ctx->launches += 1;
		RA_CUDA(cudaStreamSynchronize(ctx->stream));              // the next phase's items overwrite ctx->items / ctx->cost
	}
	if (L.cells) RA_CUDA(cudaMemcpyAsync(L.cells, &issued, sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

// Runs the chain kernel on device-resident inputs; retries once with a larger HBM ring when a part
// was wider than the rings available.
int chains_dev(ra_ctx *ctx, const float *read_dev, const float *ref_dev, const ra_anchor_t *anchors_dev,
               const ra_chain_t *chains_dev_p, const uint32_t *group_first_dev, uint32_t n_groups, int min_score_mode,
               const float *min_score_dev, float *score_dev, uint8_t *keep_dev, uint64_t *cells_host /* [2] or null */,
               uint32_t n_chains, uint64_t n_anchors, uint64_t n_read = 0, uint64_t n_ref = 0)
{
	if (n_groups == 0) { if (cells_host) cells_host[0] = cells_host[1] = 0; return RA_OK; }
	const ra_mapopt_t &o = ctx->opt;
	if (o.dtw_border_constraint != RA_DTW_BORDER_CONSTRAINT_GLOBAL && o.dtw_border_constraint != RA_DTW_BORDER_CONSTRAINT_SPARSE)
		return fail(RA_EUNSUP, "invalid border constraint (the reference exits at rmap.cpp:301-304)");
	int rc;
	if ((rc = ctx->counter.reserve(256))) return rc;
	if ((rc = ctx->cells.reserve(64))) return rc;
	const size_t warps = (size_t)ctx->n_sm * ra::kChainCtasPerSm * 4;
	uint32_t ring_cap = ctx->chain_ring.cap ? (uint32_t)(ctx->chain_ring.cap / (warps * 3 * sizeof(float))) : 0;
	{ uint32_t c = 1; while (c * 2 <= ring_cap) c *= 2; ring_cap = ring_cap ? c : 0; }
	for (int attempt = 0; attempt < 2; ++attempt) {
		ra::ChainLaunch L; memset(&L, 0, sizeof(L));
		L.read_arena = read_dev; L.ref_arena = ref_dev; L.anchors = anchors_dev; L.chains = chains_dev_p;
		L.group_first = group_first_dev; L.n_groups = n_groups;
		L.border = o.dtw_border_constraint; L.fill = o.dtw_fill_method; L.fp_contract = o.fp_contract;
		L.band_frac = o.dtw_band_radius_frac; L.match_bonus = o.dtw_match_bonus; L.min_dtw_score = o.dtw_min_score;
		L.min_score_mode = min_score_mode; L.min_score = min_score_dev; L.score = score_dev; L.keep = keep_dev;
		L.work_counter = ctx->counter.as<uint32_t>(); L.need_cap = ctx->counter.as<uint32_t>() + 4;
		L.error = ctx->counter.as<uint32_t>() + 20; L.n_read = n_read; L.n_ref = n_ref;
		L.cells = cells_host ? ctx->cells.as<uint64_t>() : nullptr;      // cell counting is instrumentation: off unless asked for
		L.ring_scratch = ring_cap ? ctx->chain_ring.as<float>() : nullptr; L.ring_cap = ring_cap;
		RA_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
		// every anchor of every chain is checked on the device before anything indexes with it (the kernels pack anchor
		// distances into narrow descriptor fields and index the arenas with them)
		RA_CUDA(cudaMemsetAsync(L.error, 0, sizeof(uint32_t), ctx->stream));
		RA_CUDA(ra::launch_validate_chains(L, 0, n_chains, ctx->stream));
		ctx->launches += 1;
		if (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_SPARSE) {
			if ((rc = sparse_pipeline(ctx, L, n_chains, n_anchors))) return rc;
		} else if (!ctx->force_generic_dtw) {
			if ((rc = global_pipeline(ctx, L, n_chains))) return rc;
		} else {
			RA_CUDA(ra::launch_chain_eval(L, ctx->n_sm, ctx->stream));
			ctx->launches += 1;
		}
		RA_CUDA(cudaEventRecord(ctx->ev[4], ctx->stream));
		ctx->timed_dtw = true;
		uint32_t need = 0, bad = 0;
		RA_CUDA(cudaMemcpyAsync(&need, L.need_cap, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(&bad, L.error, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		uint64_t cells[2] = {0, 0};
		if (L.cells) RA_CUDA(cudaMemcpyAsync(cells, L.cells, sizeof(cells), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));
		if (bad) return fail(RA_EINVAL, "a chain's anchors are not monotonic (anchors[0] must be the chain end, positions non-increasing) or leave the arenas");
		if (need == 0) { if (cells_host) { cells_host[0] = cells[0]; cells_host[1] = cells[1]; } return RA_OK; }
		if (attempt == 1) break;
		if ((rc = ctx->chain_ring.reserve(warps * 3 * (size_t)need * sizeof(float)))) return rc;
		ring_cap = need;
	}
	return fail(RA_ECUDA, "chain kernel could not obtain a wide enough ring");
}

// group_first[g] = index of the first chain of group g (groups are runs of equal ra_chain_t::group); n_groups+1 entries.
void group_boundaries(const ra_chain_t *chains, uint32_t n_chains, int min_score_mode, std::vector<uint32_t> &gf)
{
	gf.clear();
	for (uint32_t k = 0; k < n_chains; ++k)
		if (k == 0 || min_score_mode != 1 || chains[k].group != chains[k - 1].group) gf.push_back(k);
	gf.push_back(n_chains);
}

} // namespace

extern "C" {

const char *ra_last_error(void) { return g_err; }

void ra_mapopt_init(ra_mapopt_t *o)
{
	memset(o, 0, sizeof(*o));
	o->window_length1 = 3; o->window_length2 = 6;                 // roptions.c:49-50
	o->threshold1 = 4.30265f; o->threshold2 = 2.57058f;           // :51-52
	o->peak_height = 1.0f;                                        // :53
	o->dtw_border_constraint = RA_DTW_BORDER_CONSTRAINT_SPARSE;   // :37
	o->dtw_fill_method = RA_DTW_FILL_METHOD_BANDED;               // :38
	o->dtw_band_radius_frac = 0.10f;                              // :39
	o->dtw_match_bonus = 0.4f;                                    // :40
	o->dtw_min_score = 20.0f;                                     // :41
	o->fp_contract = 1;
}

int ra_init(int device, ra_ctx **out)
{
	if (!out) return fail(RA_EINVAL, "out is NULL");
	*out = nullptr;
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) { (void)cudaGetLastError(); return fail(RA_ENODEV, "no CUDA device (%s); this library has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "count=0"); }
	if (device < 0 || device >= n) return fail(RA_EINVAL, "device %d out of range [0,%d)", device, n);
	cudaDeviceProp prop;
	RA_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10) return fail(RA_ENODEV, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
	RA_CUDA(cudaSetDevice(device));
	ra_ctx *ctx = new (std::nothrow) ra_ctx();
	if (!ctx) return fail(RA_ENOMEM, "out of host memory");
	ctx->device = device; ctx->n_sm = prop.multiProcessorCount;
	ra_mapopt_init(&ctx->opt);
	e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
	if (e != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
	for (int k = 0; k < 5; ++k) {
		e = cudaEventCreate(&ctx->ev[k]);
		if (e != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaEventCreate: %s", cudaGetErrorString(e)); }
	}
	cudaStream_t *extra[3] = { &ctx->h2d, &ctx->d2h, &ctx->aux };
	for (cudaStream_t *s : extra) {
		e = cudaStreamCreateWithFlags(s, cudaStreamNonBlocking);
		if (e != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
	}
	for (int k = 0; k < kMaxSlices; ++k) {
		if ((e = cudaEventCreateWithFlags(&ctx->slice_in[k], cudaEventDisableTiming)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->slice_done[k], cudaEventDisableTiming)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->slice_dec[k], cudaEventDisableTiming)) != cudaSuccess ||
		    (e = cudaEventCreateWithFlags(&ctx->slice_scan[k], cudaEventDisableTiming)) != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaEventCreate: %s", cudaGetErrorString(e)); }
	}
	{
		int lo = 0, hi = 0;
		cudaDeviceGetStreamPriorityRange(&lo, &hi);                       // hi = numerically lowest = greatest priority
		for (int k = 0; k < kSliceStreams; ++k) {
			if ((e = cudaStreamCreateWithFlags(&ctx->slice_stream[k], cudaStreamNonBlocking)) != cudaSuccess ||
			    (e = cudaStreamCreateWithPriority(&ctx->slice_hi[k], cudaStreamNonBlocking, hi)) != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaStreamCreateWithPriority: %s", cudaGetErrorString(e)); }
			for (int j = 0; j < 4; ++j)
				if ((e = cudaEventCreateWithFlags(&ctx->split_ev[k][j], cudaEventDisableTiming)) != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaEventCreate: %s", cudaGetErrorString(e)); }
			if ((e = cudaEventCreateWithFlags(&ctx->grp_scan[k], cudaEventDisableTiming)) != cudaSuccess ||
			    (e = cudaEventCreateWithFlags(&ctx->grp_done[k], cudaEventDisableTiming)) != cudaSuccess) { delete ctx; return fail(RA_ECUDA, "cudaEventCreate: %s", cudaGetErrorString(e)); }
		}
		if (const char *g = getenv("RA_K1_GROUPS")) { const int v = atoi(g); if (v >= 1 && v <= kSliceStreams) ctx->k1_groups = v; }
		if (const char *g = getenv("RA_K1_PRIO")) ctx->k1_prio = atoi(g) != 0;
	}
	*out = ctx;
	return RA_OK;
}

void ra_destroy(ra_ctx *ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	DevBuf *bufs[] = { &ctx->ref, &ctx->sig, &ctx->chunk_off, &ctx->chunk_len, &ctx->raw, &ctx->nev, &ctx->ev_pos, &ctx->ev_off,
	                   &ctx->ev_out, &ctx->a_arena, &ctx->b_arena, &ctx->items, &ctx->cost, &ctx->counter, &ctx->ring, &ctx->order, &ctx->anchors, &ctx->chains, &ctx->group_first, &ctx->min_score, &ctx->score, &ctx->keep, &ctx->cells, &ctx->chain_ring, &ctx->sp_nparts, &ctx->sp_ntiles, &ctx->sp_pbase, &ctx->sp_tbase, &ctx->sp_tchain, &ctx->sp_cost, &ctx->sp_pcells, &ctx->sp_best, &ctx->sp_rest,
	                   &ctx->dirs, &ctx->dir_off, &ctx->slot_off, &ctx->slot_i, &ctx->slot_j, &ctx->slot_d, &ctx->path_len,
	                   &ctx->path_off, &ctx->out_i, &ctx->out_j, &ctx->out_d };
	for (DevBuf *b : bufs) b->release();
	for (int k = 0; k < 5; ++k) if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
	for (int k = 0; k < kMaxSlices; ++k) {
		if (ctx->slice_in[k]) cudaEventDestroy(ctx->slice_in[k]);
		if (ctx->slice_done[k]) cudaEventDestroy(ctx->slice_done[k]);
		if (ctx->slice_dec[k]) cudaEventDestroy(ctx->slice_dec[k]);
		if (ctx->slice_scan[k]) cudaEventDestroy(ctx->slice_scan[k]);
	}
	for (int k = 0; k < kSliceStreams; ++k) {
		if (ctx->slice_stream[k]) cudaStreamDestroy(ctx->slice_stream[k]);
		if (ctx->slice_hi[k]) cudaStreamDestroy(ctx->slice_hi[k]);
		for (int j = 0; j < 4; ++j) if (ctx->split_ev[k][j]) cudaEventDestroy(ctx->split_ev[k][j]);
		if (ctx->grp_scan[k]) cudaEventDestroy(ctx->grp_scan[k]);
		if (ctx->grp_done[k]) cudaEventDestroy(ctx->grp_done[k]);
	}
	DevBuf *dec[] = { &ctx->raw16, &ctx->rd_off, &ctx->rd_len, &ctx->rd_offset, &ctx->rd_scale, &ctx->rd_lsig };
	for (DevBuf *b : dec) b->release();
	ctx->pin_off.release(); ctx->pin_tab.release(); ctx->pin_lsig.release(); ctx->pin_gf.release();
	if (ctx->h2d) cudaStreamDestroy(ctx->h2d);
	if (ctx->d2h) cudaStreamDestroy(ctx->d2h);
	if (ctx->aux) cudaStreamDestroy(ctx->aux);
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}

int ra_set_opt(ra_ctx *ctx, const ra_mapopt_t *opt)
{
	if (!ctx || !opt) return fail(RA_EINVAL, "NULL argument");
	if (opt->dtw_border_constraint > RA_DTW_BORDER_CONSTRAINT_LOCAL) return fail(RA_EINVAL, "invalid border constraint %u", opt->dtw_border_constraint);
	if (opt->dtw_fill_method > RA_DTW_FILL_METHOD_BANDED) return fail(RA_EINVAL, "invalid fill method %u", opt->dtw_fill_method);
	ctx->opt = *opt;
	ctx->force_unfused_events = (opt->reserved & 1u) != 0;
	ctx->force_fused_events = (opt->reserved & 2u) != 0;
	ctx->force_generic_dtw = (opt->reserved & 4u) != 0;
	ctx->legacy_sparse_tiles = (opt->reserved & 8u) != 0;
	return RA_OK;
}

uint64_t ra_launch_count(const ra_ctx *ctx) { return ctx ? ctx->launches : 0; }
int ra_get_opt(const ra_ctx *ctx, ra_mapopt_t *opt) { if (!ctx || !opt) return fail(RA_EINVAL, "NULL argument"); *opt = ctx->opt; return RA_OK; }
int ra_set_tuning(ra_ctx *ctx, const char *name, int64_t value)
{
	if (!ctx || !name) return fail(RA_EINVAL, "NULL argument");
	if (!strcmp(name, "chain_slice_min_anchors")) { if (value < 1) return fail(RA_EINVAL, "chain_slice_min_anchors must be >= 1"); ctx->chain_slice_min_anchors = (uint64_t)value; return RA_OK; }
	if (!strcmp(name, "k1_groups")) { if (value < 1 || value > kSliceStreams) return fail(RA_EINVAL, "k1_groups must be 1..%d", kSliceStreams); ctx->k1_groups = (int)value; return RA_OK; }
	if (!strcmp(name, "k1_priority_stream")) { ctx->k1_prio = value != 0; return RA_OK; }
	if (!strcmp(name, "map_sig_whole")) { ctx->map_sig_whole = value != 0; return RA_OK; }
	if (!strcmp(name, "map_host_gather")) { if (value < 0 || value > 16) return fail(RA_EINVAL, "map_host_gather must be 0..16 (threads)"); ctx->map_host_gather = (int)value; return RA_OK; }
	return fail(RA_EINVAL, "unknown tuning parameter %s", name);
}
// internal (map.cu): kernels launched outside this file count too; errors raised there go through ra_last_error()
void ra__add_launches(ra_ctx *ctx, uint64_t n) { if (ctx) ctx->launches += n; }
int ra__device(const ra_ctx *ctx) { return ctx ? ctx->device : 0; }
int ra__map_sig_whole(const ra_ctx *ctx) { return ctx && ctx->map_sig_whole; }
int ra__map_host_gather(const ra_ctx *ctx) { return ctx ? ctx->map_host_gather : 0; }
// internal (map.cu): lets the entry points that read "the events resident on the device" read the map loop's event arena
void ra__swap_resident(ra_ctx *ctx, void **dev_ptr, uint64_t *n_floats)
{
	std::swap(ctx->ev_out.p, *dev_ptr);
	std::swap(ctx->resident_events, *n_floats);
}
int ra__fail(int code, const char *fmt, ...)
{
	va_list ap; va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
	return code;
}

int ra_last_timing(ra_ctx *ctx, float *events_scan_ms, float *events_total_ms, float *dtw_ms)
{
	if (!ctx) return fail(RA_EINVAL, "NULL ctx");
	RA_CUDA(cudaSetDevice(ctx->device));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	float a = 0.f, b = 0.f, c = 0.f;
	if (ctx->timed_k1) { RA_CUDA(cudaEventElapsedTime(&a, ctx->ev[0], ctx->ev[1])); RA_CUDA(cudaEventElapsedTime(&b, ctx->ev[0], ctx->ev[2])); }
	if (ctx->timed_dtw) RA_CUDA(cudaEventElapsedTime(&c, ctx->ev[3], ctx->ev[4]));
	if (events_scan_ms) *events_scan_ms = a;
	if (events_total_ms) *events_total_ms = b;
	if (dtw_ms) *dtw_ms = c;
	return RA_OK;
}

/* ---- reference mirror ---------------------------------------------------------------------------- */

int ra_ref_alloc(ra_ctx *ctx, uint64_t n_floats)
{
	if (!ctx) return fail(RA_EINVAL, "NULL ctx");
	RA_CUDA(cudaSetDevice(ctx->device));
	int rc = ctx->ref.reserve((size_t)std::max<uint64_t>(n_floats, 1) * sizeof(float));
	if (rc) return rc;
	ctx->ref_n = n_floats;
	return RA_OK;
}

int ra_ref_load_flat(ra_ctx *ctx, const float *ref, uint64_t n_floats)
{
	if (!ctx || (!ref && n_floats)) return fail(RA_EINVAL, "NULL argument");
	int rc = ra_ref_alloc(ctx, n_floats);
	if (rc) return rc;
	RA_CUDA(cudaMemcpyAsync(ctx->ref.p, ref, (size_t)n_floats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

int ra_ref_load(ra_ctx *ctx, uint32_t n_seq, const float *const *fwd, const float *const *rev, const uint32_t *len,
                uint64_t *arena_off)
{
	if (!ctx || !fwd || !rev || !len || !arena_off) return fail(RA_EINVAL, "NULL argument");
	uint64_t tot = 0;
	for (uint32_t s = 0; s < n_seq; ++s) { arena_off[2 * s] = tot; tot += len[s]; arena_off[2 * s + 1] = tot; tot += len[s]; }
	int rc = ra_ref_alloc(ctx, tot);
	if (rc) return rc;
	for (uint32_t s = 0; s < n_seq; ++s) {
		RA_CUDA(cudaMemcpyAsync(ctx->ref.as<float>() + arena_off[2 * s], fwd[s], (size_t)len[s] * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(ctx->ref.as<float>() + arena_off[2 * s + 1], rev[s], (size_t)len[s] * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	}
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

int ra_ref_device(ra_ctx *ctx, float **dev_ptr, uint64_t *n_floats)
{
	if (!ctx || !dev_ptr || !n_floats) return fail(RA_EINVAL, "NULL argument");
	*dev_ptr = ctx->ref.as<float>(); *n_floats = ctx->ref_n;
	return RA_OK;
}

/* ---- K1 ------------------------------------------------------------------------------------------ */

int ra_detect_events_batch_dev(ra_ctx *ctx, const float *sig_dev, const uint64_t *chunk_off_dev,
                               const uint32_t *chunk_len_dev, uint32_t n_chunks, uint32_t max_chunk_len,
                               float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev, uint64_t *total_events)
{
	if (!ctx || !ev_off_dev || (n_chunks && (!sig_dev || !chunk_off_dev || !chunk_len_dev || !ev_out_dev)))
		return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	return events_dev(ctx, sig_dev, chunk_off_dev, chunk_len_dev, n_chunks, max_chunk_len, ev_out_dev, ev_cap, ev_off_dev, total_events);
}

int ra_detect_events_batch(ra_ctx *ctx, const float *sig, uint64_t n_sig, const uint64_t *chunk_off,
                           const uint32_t *chunk_len, uint32_t n_chunks, float *ev_out, uint64_t ev_cap, uint64_t *ev_off)
{
	if (!ctx || !ev_off || (n_chunks && (!sig || !chunk_off || !chunk_len || !ev_out))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	uint32_t max_len = 0; uint64_t sum_len = 0;
	for (uint32_t c = 0; c < n_chunks; ++c) {
		if (chunk_off[c] + chunk_len[c] > n_sig) return fail(RA_EINVAL, "chunk %u exceeds the signal array", c);
		max_len = std::max(max_len, chunk_len[c]); sum_len += chunk_len[c];
	}
	RA_CUDA(cudaStreamSynchronize(ctx->stream));                   // the staging buffers below may be reallocated
	int rc;
	if ((rc = ctx->sig.reserve((size_t)std::max<uint64_t>(n_sig, 4) * sizeof(float)))) return rc;
	if ((rc = ctx->chunk_off.reserve((size_t)(n_chunks + 1) * sizeof(uint64_t)))) return rc;
	if ((rc = ctx->chunk_len.reserve((size_t)(n_chunks + 1) * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->ev_off.reserve((size_t)(n_chunks + 1) * sizeof(uint64_t)))) return rc;
	// the same bound events_pipe_layout sizes the peak lists with: emissions of one detector are >= w/2 + 2 samples apart
	const uint64_t dev_cap = std::min<uint64_t>(ev_cap, sum_len / (ctx->opt.window_length1 / 2 + 2) + sum_len / (ctx->opt.window_length2 / 2 + 2) + 4ull * n_chunks + 16);
	if ((rc = ctx->ev_out.reserve((size_t)std::max<uint64_t>(dev_cap, 1) * sizeof(float)))) return rc;
	if (n_chunks == 0) { ev_off[0] = 0; ctx->last_events = ctx->resident_events = 0; return RA_OK; }
	// Slices of chunks; slice k may start once sig[0 .. need[k]) is on the device (need = running max of chunk ends).
	const int n_slices = (int)std::min<uint64_t>(kMaxSlices, std::max<uint64_t>(1, sum_len / (8u << 20)));   // >= 32 MB of signal per slice
	uint32_t cut[kMaxSlices + 1]; uint64_t need[kMaxSlices + 1];
	cut[0] = 0; need[0] = 0;
	{
		uint64_t run_max = 0; uint32_t c = 0;
		for (int k = 1; k <= n_slices; ++k) {
			const uint32_t end = (uint32_t)((uint64_t)n_chunks * k / n_slices);
			for (; c < end; ++c) run_max = std::max<uint64_t>(run_max, chunk_off[c] + chunk_len[c]);
			cut[k] = end; need[k] = (k == n_slices) ? std::max(run_max, need[k - 1]) : run_max;
		}
	}
	RA_CUDA(cudaMemcpyAsync(ctx->chunk_off.p, chunk_off, (size_t)n_chunks * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->h2d));
	RA_CUDA(cudaMemcpyAsync(ctx->chunk_len.p, chunk_len, (size_t)n_chunks * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->h2d));
	for (int k = 0; k < n_slices; ++k) {
		if (need[k + 1] > need[k])
			RA_CUDA(cudaMemcpyAsync(ctx->sig.as<float>() + need[k], sig + need[k], (size_t)(need[k + 1] - need[k]) * sizeof(float), cudaMemcpyHostToDevice, ctx->h2d));
		RA_CUDA(cudaEventRecord(ctx->slice_in[k], ctx->h2d));
	}
	if ((rc = ctx->pin_off.reserve((size_t)(n_chunks + 1) * sizeof(uint64_t)))) return rc;
	if ((rc = slices_prepare(ctx, n_chunks, max_len, dev_cap))) return rc;
	for (int k = 0; k < n_slices; ++k)
		if ((rc = slice_launch(ctx, ctx->sig.as<float>(), k, cut[k], cut[k + 1] - cut[k], max_len, dev_cap))) return rc;
	return slices_collect(ctx, cut, n_slices, dev_cap, ev_out, ev_cap, ev_off, n_chunks);
}

/* Raw-signal form: int16 DAC samples in, decoded and filtered on the device exactly as
 * ri_read_sig does (rsig.cpp:213-222), cut into chunks as map_worker_for does (rmap.cpp:684-689). */
int ra_detect_events_raw_batch(ra_ctx *ctx, const int16_t *raw, uint64_t n_raw, const uint64_t *raw_off, const uint32_t *raw_len,
                               const float *offset, const float *scale, uint32_t n_reads, uint32_t chunk_size, uint32_t max_chunks,
                               uint32_t *l_sig, uint32_t *chunk_first, float *ev_out, uint64_t ev_cap, uint64_t *ev_off)
{
	if (!ctx || !ev_off || !chunk_first || !l_sig || (n_reads && (!raw || !raw_off || !raw_len || !offset || !scale || !ev_out))) return fail(RA_EINVAL, "NULL argument");
	if (chunk_size == 0 || max_chunks == 0) return fail(RA_EINVAL, "chunk_size and max_chunks must be positive");
	RA_CUDA(cudaSetDevice(ctx->device));
	uint64_t max_chunks_total = 0, max_samples = 0;
	for (uint32_t r = 0; r < n_reads; ++r) {
		if (raw_off[r] + raw_len[r] > n_raw) return fail(RA_EINVAL, "read %u exceeds the raw array", r);
		if (r && raw_off[r] < raw_off[r - 1] + raw_len[r - 1]) return fail(RA_EINVAL, "reads must be laid out in order without overlap (read %u)", r);
		const uint64_t nch = std::min<uint64_t>(max_chunks, ((uint64_t)raw_len[r] + chunk_size - 1) / chunk_size);
		max_chunks_total += nch; max_samples += std::min<uint64_t>(raw_len[r], nch * chunk_size);
	}
	if (max_chunks_total >= 0xffffffffull) return fail(RA_EINVAL, "too many chunks in one batch");
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	int rc;
	if ((rc = ctx->raw16.reserve((size_t)std::max<uint64_t>(n_raw, 8) * sizeof(int16_t) + 16))) return rc;
	if ((rc = ctx->sig.reserve((size_t)std::max<uint64_t>(n_raw, 4) * sizeof(float)))) return rc;
	if ((rc = ctx->rd_off.reserve((size_t)(n_reads + 1) * 8))) return rc;
	if ((rc = ctx->rd_len.reserve((size_t)(n_reads + 1) * 4))) return rc;
	if ((rc = ctx->rd_offset.reserve((size_t)(n_reads + 1) * 4))) return rc;
	if ((rc = ctx->rd_scale.reserve((size_t)(n_reads + 1) * 4))) return rc;
	if ((rc = ctx->rd_lsig.reserve((size_t)(n_reads + 1) * 4))) return rc;
	if ((rc = ctx->chunk_off.reserve((size_t)(max_chunks_total + 1) * sizeof(uint64_t)))) return rc;
	if ((rc = ctx->chunk_len.reserve((size_t)(max_chunks_total + 1) * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->ev_off.reserve((size_t)(max_chunks_total + 1) * sizeof(uint64_t)))) return rc;
	const uint64_t dev_cap = std::min<uint64_t>(ev_cap, max_samples / (ctx->opt.window_length1 / 2 + 2) + max_samples / (ctx->opt.window_length2 / 2 + 2) + 4ull * max_chunks_total + 16);
	if ((rc = ctx->ev_out.reserve((size_t)std::max<uint64_t>(dev_cap, 1) * sizeof(float)))) return rc;
	if ((rc = ctx->pin_lsig.reserve((size_t)(n_reads + 1) * 4))) return rc;
	chunk_first[0] = 0;
	if (n_reads == 0) { ev_off[0] = 0; ctx->last_events = ctx->resident_events = 0; return RA_OK; }
	// read slices of roughly equal raw size
	const int n_slices = (int)std::min<uint64_t>(kMaxSlices, std::max<uint64_t>(1, n_raw / (16u << 20)));     // >= 32 MB of int16 per slice
	uint32_t rcut[kMaxSlices + 1]; rcut[0] = 0;
	{
		uint32_t r = 0;
		for (int k = 1; k <= n_slices; ++k) {
			const uint64_t target = (k == n_slices) ? n_raw : n_raw * k / n_slices;
			while (r < n_reads && raw_off[r] + raw_len[r] <= target) ++r;
			if (k == n_slices) r = n_reads;
			rcut[k] = r;
		}
	}
	Trace tr; tr.mark("start");
	if (tr.on) { trace_events_init(); cudaEventRecord(g_tr_t0, ctx->h2d); }
	RA_CUDA(cudaMemcpyAsync(ctx->rd_off.p, raw_off, (size_t)n_reads * 8, cudaMemcpyHostToDevice, ctx->h2d));
	RA_CUDA(cudaMemcpyAsync(ctx->rd_len.p, raw_len, (size_t)n_reads * 4, cudaMemcpyHostToDevice, ctx->h2d));
	RA_CUDA(cudaMemcpyAsync(ctx->rd_offset.p, offset, (size_t)n_reads * 4, cudaMemcpyHostToDevice, ctx->h2d));
	RA_CUDA(cudaMemcpyAsync(ctx->rd_scale.p, scale, (size_t)n_reads * 4, cudaMemcpyHostToDevice, ctx->h2d));
	for (int k = 0; k < n_slices; ++k) {
		if (rcut[k + 1] > rcut[k]) {
			const uint64_t b0 = raw_off[rcut[k]], b1 = raw_off[rcut[k + 1] - 1] + raw_len[rcut[k + 1] - 1];
			RA_CUDA(cudaMemcpyAsync(ctx->raw16.as<int16_t>() + b0, raw + b0, (size_t)(b1 - b0) * sizeof(int16_t), cudaMemcpyHostToDevice, ctx->h2d));
		}
		RA_CUDA(cudaEventRecord(ctx->slice_in[k], ctx->h2d));
	}
	// decode every slice on the aux stream as it arrives; l_sig comes back per slice (the host cuts the chunks)
	for (int k = 0; k < n_slices; ++k) {
		const uint32_t r0 = rcut[k], nr = rcut[k + 1] - r0;
		RA_CUDA(cudaStreamWaitEvent(ctx->aux, ctx->slice_in[k], 0));
		RA_CUDA(ra::launch_decode_raw(ctx->raw16.as<int16_t>(), ctx->rd_off.as<uint64_t>() + r0, ctx->rd_len.as<uint32_t>() + r0, ctx->rd_offset.as<float>() + r0,
		                              ctx->rd_scale.as<float>() + r0, nr, ctx->sig.as<float>(), ctx->rd_lsig.as<uint32_t>() + r0, ctx->aux));
		if (nr) { ctx->launches += 1; RA_CUDA(cudaMemcpyAsync(ctx->pin_lsig.as<uint32_t>() + r0, ctx->rd_lsig.as<uint32_t>() + r0, (size_t)nr * 4, cudaMemcpyDeviceToHost, ctx->aux)); }
		RA_CUDA(cudaEventRecord(ctx->slice_dec[k], ctx->aux));
	}
	if ((rc = ctx->pin_off.reserve((size_t)(max_chunks_total + 1) * sizeof(uint64_t)))) return rc;
	if ((rc = ctx->pin_tab.reserve((size_t)(max_chunks_total + 1) * 12))) return rc;
	if ((rc = slices_prepare(ctx, (uint32_t)max_chunks_total, chunk_size, dev_cap))) return rc;
	uint64_t *coff = ctx->pin_tab.as<uint64_t>();                               // pinned, so the table uploads are asynchronous
	uint32_t *clen = reinterpret_cast<uint32_t *>(coff + max_chunks_total + 1);
	uint32_t cut[kMaxSlices + 1]; cut[0] = 0;
	uint32_t n_chunks = 0;
	const uint32_t *lsig_stage = ctx->pin_lsig.as<uint32_t>();
	for (int k = 0; k < n_slices; ++k) {
		RA_CUDA(cudaEventSynchronize(ctx->slice_dec[k]));
		tr.mark("decoded", k);
		const uint32_t c_begin = n_chunks;
		for (uint32_t r = rcut[k]; r < rcut[k + 1]; ++r) {
			const uint32_t qlen = lsig_stage[r];
			l_sig[r] = qlen;
			uint32_t cs = 0, cc = 0;
			for (; cs < qlen && cc < max_chunks; cs += chunk_size, ++cc) {       // rmap.cpp:684-689
				coff[n_chunks] = raw_off[r] + cs; clen[n_chunks] = std::min(chunk_size, qlen - cs); ++n_chunks;
			}
			chunk_first[r + 1] = n_chunks;
		}
		const uint32_t n = n_chunks - c_begin;
		cut[k + 1] = n_chunks;
		// the table goes up on the stream that runs this slice's K1 (the aux stream still holds the decodes of later slices,
		// which wait for their copies: anything queued behind them would start late)
		cudaStream_t ks = ctx->slices_concurrent ? ctx->slice_stream[k % kCopySliceStreams] : ctx->stream;
		if (n) {
			RA_CUDA(ra::launch_fetch_tables(coff + c_begin, clen + c_begin, ctx->chunk_off.as<uint64_t>() + c_begin, ctx->chunk_len.as<uint32_t>() + c_begin, n, ks));
			ctx->launches += 1;
		}
		RA_CUDA(cudaEventRecord(ctx->slice_in[k], ks));                         // slice k is ready for K1: signal decoded (host saw slice_dec[k]), table on the device
		if ((rc = slice_launch(ctx, ctx->sig.as<float>(), k, c_begin, n, chunk_size, dev_cap))) return rc;
	}
	tr.mark("all launched");
	rc = slices_collect(ctx, cut, n_slices, dev_cap, ev_out, ev_cap, ev_off, n_chunks);
	RA_CUDA(cudaStreamSynchronize(ctx->aux));
	tr.mark("collected");
	if (tr.on && ctx->slices_concurrent)
		for (int k = 0; k < n_slices; ++k) {
			float a0 = 0, a1 = 0, a2 = 0;
			cudaEventElapsedTime(&a0, g_tr_t0, g_tr_ev[k][0]); cudaEventElapsedTime(&a1, g_tr_t0, g_tr_ev[k][1]); cudaEventElapsedTime(&a2, g_tr_t0, g_tr_ev[k][2]);
			fprintf(stderr, "[ra trace] device: slice %d K1 start %.3f  K1 end %.3f  offsets copied %.3f ms\n", k, a0, a1, a2);
		}
	return rc;
}

int ra_last_event_positions(ra_ctx *ctx, uint32_t *pos_out, uint64_t cap)
{
	if (!ctx || !pos_out) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	if (cap < ctx->last_events) return fail(RA_EINVAL, "capacity %llu < %llu events", (unsigned long long)cap, (unsigned long long)ctx->last_events);
	if (ctx->last_events) RA_CUDA(cudaMemcpyAsync(pos_out, ctx->ev_pos.p, (size_t)ctx->last_events * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

/* ---- K2 / K3 ------------------------------------------------------------------------------------- */

uint64_t ra_dtw_cells(uint32_t a_len, uint32_t b_len, int32_t band_radius)
{
	if (band_radius < 0) return (uint64_t)a_len * b_len;
	if (a_len == 0 || b_len == 0) return 0;
	const int64_t n = std::max(a_len, b_len), m = std::min(a_len, b_len);
	const uint32_t extra = (((uint32_t)n - (uint32_t)m) * (uint32_t)band_radius + (uint32_t)n - 1u) / (uint32_t)n;   // dtw.cpp:298
	const int64_t R = (int64_t)band_radius + (int64_t)extra;
	const int64_t P = R + (R % 2 == 0 ? 1 : 0), S = R + (R % 2 == 1 ? 1 : 0);
	auto count = [&](int64_t start_i, int64_t start_j, int64_t len) -> uint64_t {
		int64_t o0 = std::max<int64_t>(std::max<int64_t>(0, start_i - (n - 1)), -start_j);                              // :365, :419
		int64_t o1 = std::min<int64_t>(std::min<int64_t>(len, start_i + 1), m - start_j);                               // :366, :420
		return o1 > o0 ? (uint64_t)(o1 - o0) : 0;
	};
	uint64_t cells = 1;
	int64_t r = 0;
	for (int64_t c = 1; c < n; ++c) {
		if ((r + 1) * n <= m * c) { ++r; cells += count(c + S / 2 - 1, r - S / 2, S); }
		cells += count(c + P / 2, r - P / 2, P);
	}
	return cells;
}

int ra_dtw_batch_dev(ra_ctx *ctx, const float *a_arena_dev, const float *b_arena_dev, const ra_dtw_item_t *items_dev,
                     const ra_dtw_item_t *items_host, uint32_t n_items, float *cost_out_dev)
{
	if (!ctx || (n_items && (!a_arena_dev || !items_dev || !items_host || !cost_out_dev))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	const float *b = b_arena_dev ? b_arena_dev : ctx->ref.as<float>();
	if (n_items && !b) return fail(RA_EINVAL, "no reference arena: pass b_arena or call ra_ref_load first");
	int rc = dtw_dev(ctx, a_arena_dev, b, items_dev, items_host, n_items, cost_out_dev);
	if (rc) return rc;
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

static int check_items(const ra_dtw_item_t *items, uint32_t n_items, uint64_t n_a, uint64_t n_b)
{
	for (uint32_t k = 0; k < n_items; ++k) {
		const ra_dtw_item_t &it = items[k];
		if (it.a_len == 0 || it.b_len == 0) return fail(RA_EINVAL, "item %u has an empty span (the reference asserts a_length > 0 && b_length > 0)", k);
		if (it.a_off + it.a_len > n_a) return fail(RA_EINVAL, "item %u: read span exceeds the read arena", k);
		if (it.b_off + it.b_len > n_b) return fail(RA_EINVAL, "item %u: reference span exceeds the reference arena", k);
		if (it.a_len > 0x3fffffffu || it.b_len > 0x3fffffffu) return fail(RA_EINVAL, "item %u: span too long", k);
	}
	return RA_OK;
}

int ra_dtw_batch(ra_ctx *ctx, const float *a_arena, uint64_t n_a, const float *b_arena, uint64_t n_b,
                 const ra_dtw_item_t *items, uint32_t n_items, float *cost_out)
{
	if (!ctx || (n_items && (!a_arena || !items || !cost_out))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	if (n_items == 0) return RA_OK;
	if (!b_arena && !ctx->ref.p) return fail(RA_EINVAL, "no reference arena: pass b_arena or call ra_ref_load first");
	int rc = check_items(items, n_items, n_a, b_arena ? n_b : ctx->ref_n);
	if (rc) return rc;
	if ((rc = ctx->a_arena.reserve((size_t)n_a * sizeof(float)))) return rc;
	if ((rc = ctx->items.reserve((size_t)n_items * sizeof(ra_dtw_item_t)))) return rc;
	if ((rc = ctx->cost.reserve((size_t)n_items * sizeof(float)))) return rc;
	RA_CUDA(cudaMemcpyAsync(ctx->a_arena.p, a_arena, (size_t)n_a * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	const float *b_dev = ctx->ref.as<float>();
	if (b_arena) {
		if ((rc = ctx->b_arena.reserve((size_t)n_b * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->b_arena.p, b_arena, (size_t)n_b * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		b_dev = ctx->b_arena.as<float>();
	}
	RA_CUDA(cudaMemcpyAsync(ctx->items.p, items, (size_t)n_items * sizeof(ra_dtw_item_t), cudaMemcpyHostToDevice, ctx->stream));
	if ((rc = dtw_dev(ctx, ctx->a_arena.as<float>(), b_dev, ctx->items.as<ra_dtw_item_t>(), items, n_items, ctx->cost.as<float>()))) return rc;
	RA_CUDA(cudaMemcpyAsync(cost_out, ctx->cost.p, (size_t)n_items * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	return RA_OK;
}

/* ---- K4 ------------------------------------------------------------------------------------------ */

int ra_dtw_tb_batch(ra_ctx *ctx, const float *a_arena, uint64_t n_a, const float *b_arena, uint64_t n_b,
                    const ra_dtw_item_t *items, uint32_t n_items, float *cost_out, uint32_t *path_i, uint32_t *path_j,
                    float *path_diff, uint64_t path_cap, uint64_t *path_off)
{
	if (!ctx || !path_off || (n_items && (!items || !cost_out || !path_i || !path_j))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	path_off[0] = 0;
	if (n_items == 0) return RA_OK;
	if (!a_arena && !ctx->resident_events) return fail(RA_EINVAL, "a_arena is NULL and no events are resident on the device");
	if (!a_arena) n_a = ctx->resident_events;                     // a_arena == NULL: the events the last event-detection call left on the device
	if (!b_arena && !ctx->ref.p) return fail(RA_EINVAL, "no reference arena: pass b_arena or call ra_ref_load first");
	int rc = check_items(items, n_items, n_a, b_arena ? n_b : ctx->ref_n);
	if (rc) return rc;
	// per-item direction matrix (diagonal-major, 16 cells per word, 128-byte aligned) and path slot
	std::vector<uint64_t> dir_off(n_items + 1), slot_off(n_items + 1);
	uint64_t dwords = 0, slots = 0; uint32_t maxw = 0;
	for (uint32_t k = 0; k < n_items; ++k) {
		const ra_dtw_item_t &it = items[k];
		dir_off[k] = dwords; slot_off[k] = slots;
		const uint64_t stride = ((uint64_t)it.b_len + 15) >> 4;
		dwords += (((uint64_t)it.a_len + it.b_len - 1) * stride + 31) & ~(uint64_t)31;
		slots += (uint64_t)it.a_len + it.b_len - 1;
		maxw = std::max(maxw, it.b_len);
	}
	dir_off[n_items] = dwords; slot_off[n_items] = slots;
	if (a_arena && (rc = ctx->a_arena.reserve((size_t)n_a * sizeof(float)))) return rc;
	if ((rc = ctx->items.reserve((size_t)n_items * sizeof(ra_dtw_item_t)))) return rc;
	if ((rc = ctx->cost.reserve((size_t)n_items * sizeof(float)))) return rc;
	if ((rc = ctx->dirs.reserve((size_t)dwords * 4))) return rc;
	if ((rc = ctx->dir_off.reserve((size_t)(n_items + 1) * 8))) return rc;
	if ((rc = ctx->slot_off.reserve((size_t)(n_items + 1) * 8))) return rc;
	if ((rc = ctx->slot_i.reserve((size_t)slots * 4))) return rc;
	if ((rc = ctx->slot_j.reserve((size_t)slots * 4))) return rc;
	if (path_diff && (rc = ctx->slot_d.reserve((size_t)slots * 4))) return rc;
	if ((rc = ctx->path_len.reserve((size_t)n_items * 4))) return rc;
	if ((rc = ctx->path_off.reserve((size_t)(n_items + 1) * 8))) return rc;
	if ((rc = ctx->out_i.reserve((size_t)slots * 4))) return rc;
	if ((rc = ctx->out_j.reserve((size_t)slots * 4))) return rc;
	if (path_diff && (rc = ctx->out_d.reserve((size_t)slots * 4))) return rc;
	if ((rc = ctx->counter.reserve(256))) return rc;
	if (a_arena) RA_CUDA(cudaMemcpyAsync(ctx->a_arena.p, a_arena, (size_t)n_a * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	const float *a_dev = a_arena ? ctx->a_arena.as<float>() : ctx->ev_out.as<float>();
	const float *b_dev = ctx->ref.as<float>();
	if (b_arena) {
		if ((rc = ctx->b_arena.reserve((size_t)n_b * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->b_arena.p, b_arena, (size_t)n_b * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		b_dev = ctx->b_arena.as<float>();
	}
	RA_CUDA(cudaMemcpyAsync(ctx->items.p, items, (size_t)n_items * sizeof(ra_dtw_item_t), cudaMemcpyHostToDevice, ctx->stream));
	RA_CUDA(cudaMemcpyAsync(ctx->dir_off.p, dir_off.data(), (size_t)(n_items + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
	RA_CUDA(cudaMemcpyAsync(ctx->slot_off.p, slot_off.data(), (size_t)(n_items + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));

	ra::DtwLaunch L; memset(&L, 0, sizeof(L));
	L.a_arena = a_dev; L.b_arena = b_dev; L.items = ctx->items.as<ra_dtw_item_t>(); L.n_items = n_items;
	L.cost = ctx->cost.as<float>(); L.work_counter = ctx->counter.as<uint32_t>();
	if (maxw > ra::dtw_generic_smem_ring_cap()) {
		uint32_t cap = 1; while (cap < maxw) cap <<= 1;
		const size_t warps = (size_t)ctx->n_sm * 4 * 4;
		if ((rc = ctx->ring.reserve(warps * 3 * cap * sizeof(float)))) return rc;
		L.ring_scratch = ctx->ring.as<float>(); L.ring_cap = cap;
	}
	L.dirs = ctx->dirs.as<uint32_t>(); L.dir_off = ctx->dir_off.as<uint64_t>();
	L.path_i = ctx->slot_i.as<uint32_t>(); L.path_j = ctx->slot_j.as<uint32_t>();
	L.path_diff = path_diff ? ctx->slot_d.as<float>() : nullptr;
	L.slot_off = ctx->slot_off.as<uint64_t>(); L.path_len = ctx->path_len.as<uint32_t>();
	int nl = 0;
	RA_CUDA(ra::launch_dtw_generic(L, true, ctx->n_sm, ctx->stream, &nl));
	RA_CUDA(ra::launch_exclusive_scan_u32(ctx->path_len.as<uint32_t>(), ctx->path_off.as<uint64_t>(), n_items, ctx->stream));
	RA_CUDA(ra::launch_path_compact(ctx->slot_i.as<uint32_t>(), ctx->slot_j.as<uint32_t>(), L.path_diff, ctx->slot_off.as<uint64_t>(),
	                                ctx->items.as<ra_dtw_item_t>(), ctx->path_len.as<uint32_t>(), ctx->path_off.as<uint64_t>(), n_items,
	                                ctx->out_i.as<uint32_t>(), ctx->out_j.as<uint32_t>(), path_diff ? ctx->out_d.as<float>() : nullptr, ctx->stream));
	ctx->launches += nl + 2;
	RA_CUDA(cudaMemcpyAsync(path_off, ctx->path_off.p, (size_t)(n_items + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaMemcpyAsync(cost_out, ctx->cost.p, (size_t)n_items * 4, cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	const uint64_t total = path_off[n_items];
	if (total > path_cap) return fail(RA_EINVAL, "path capacity %llu < %llu", (unsigned long long)path_cap, (unsigned long long)total);
	if (total) {
		RA_CUDA(cudaMemcpyAsync(path_i, ctx->out_i.p, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(path_j, ctx->out_j.p, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
		if (path_diff) RA_CUDA(cudaMemcpyAsync(path_diff, ctx->out_d.p, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));
	}
	return RA_OK;
}

/* ---- chain level --------------------------------------------------------------------------------- */

// The sub-problems align_chain would issue for one chain (rmap.cpp:192-300), in the reference's order.
static int chain_items(const ra_mapopt_t &o, const ra_anchor_t *anc, const ra_chain_t &ch, bool cigar,
                       std::vector<ra_dtw_item_t> &items)
{
	if (ch.n_anchors == 0) return fail(RA_EINVAL, "chain without anchors");
	const bool banded = (o.dtw_fill_method == RA_DTW_FILL_METHOD_BANDED);
	for (uint32_t p = 1; p < ch.n_anchors; ++p)                    // anchors[0] is the chain end; positions never increase along the array
		if (anc[p].query_position > anc[p - 1].query_position || anc[p].target_position > anc[p - 1].target_position)
			return fail(RA_EINVAL, "chain anchors are not monotonic (anchor %u)", p);
	auto push = [&](const ra_anchor_t &s, const ra_anchor_t &e, bool excl) {
		ra_dtw_item_t it;
		it.a_off = ch.read_off + s.query_position;                  // rmap.cpp:201, :259
		it.b_off = ch.ref_off + s.target_position;                  // :198, :256
		it.a_len = e.query_position - s.query_position + 1;         // :202, :260
		it.b_len = e.target_position - s.target_position + 1;       // :199, :257
		if (cigar || !banded) it.band_radius = -1;                  // DTW_global / DTW_global_tb
		else it.band_radius = std::max(1, (int)((float)it.a_len * o.dtw_band_radius_frac));   // :214, :276
		it.exclude_last = excl ? 1u : 0u;
		items.push_back(it);
	};
	if (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_GLOBAL) {
		if (cigar && banded) return fail(RA_EUNSUP, "global border + banded fill + cigar is not implemented by the reference (rmap.cpp:223-225 asserts)");
		push(anc[ch.n_anchors - 1], anc[0], false);
	} else if (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_SPARSE) {
		const uint32_t parts = ch.n_anchors - 1;
		for (uint32_t p = 0; p < parts; ++p)
			push(anc[parts - p], anc[parts - p - 1], cigar ? false : (p != parts - 1));     // :270; cigar ignores it (:283-284)
	} else {
		return fail(RA_EUNSUP, "invalid border constraint (the reference exits at rmap.cpp:301-304)");
	}
	return RA_OK;
}

static inline float chain_score(const ra_mapopt_t &o, uint32_t n_aligned, float cost)
{
	// rmap.cpp:306; the reference build (-O3, -ffp-contract=fast default of GNU C++) fuses it
	if (o.fp_contract) return fmaf((float)n_aligned, o.dtw_match_bonus, -cost);
	volatile float prod = (float)n_aligned * o.dtw_match_bonus;
	return prod - cost;
}

int ra_align_chains_batch(ra_ctx *ctx, const float *read_arena, uint64_t n_read, const float *ref_arena, uint64_t n_ref,
                          const ra_anchor_t *anchors, uint64_t n_anchors, const ra_chain_t *chains, uint32_t n_chains,
                          int min_score_mode, const float *min_score, float *alignment_score, uint8_t *keep,
                          uint64_t *cells_issued, uint64_t *cells_useful)
{
	if (!ctx || (n_chains && (!anchors || !chains || !alignment_score))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	Trace tr; tr.mark("chains: start");
	if (n_chains == 0) { if (cells_issued) *cells_issued = 0; if (cells_useful) *cells_useful = 0; return RA_OK; }
	if (!read_arena && !ctx->resident_events) return fail(RA_EINVAL, "read_arena is NULL and no events are resident on the device");
	if (!ref_arena && !ctx->ref.p) return fail(RA_EINVAL, "no reference arena: pass ref_arena or call ra_ref_load first");
	const uint64_t nr = read_arena ? n_read : ctx->resident_events, nf = ref_arena ? n_ref : ctx->ref_n;
	// bounds the kernels trust, pass 1: the chain records only (sequential, cheap); pass 2, which touches the anchors, runs
	// after the uploads have been queued so that it overlaps them
	bool laid_in_order = true;
	for (uint32_t k = 0; k < n_chains; ++k) {
		const ra_chain_t &ch = chains[k];
		if (ch.n_anchors < 1 || ch.anchor_off + ch.n_anchors > n_anchors) return fail(RA_EINVAL, "chain %u: anchors out of range", k);
		if (k && ch.anchor_off < chains[k - 1].anchor_off + chains[k - 1].n_anchors) laid_in_order = false;
	}
	auto validate_spans = [&](uint32_t kb, uint32_t ke) -> int {
		for (uint32_t k = kb; k < ke; ++k) {
			const ra_chain_t &ch = chains[k];
			const ra_anchor_t &s = anchors[ch.anchor_off + ch.n_anchors - 1], &e = anchors[ch.anchor_off];
			if (e.query_position < s.query_position || e.target_position < s.target_position) return fail(RA_EINVAL, "chain %u: anchors[0] must be the chain end", k);
			if (ch.read_off + e.query_position >= nr) return fail(RA_EINVAL, "chain %u: read span exceeds the read arena", k);
			if (ch.ref_off + e.target_position >= nf) return fail(RA_EINVAL, "chain %u: reference span exceeds the reference arena", k);
		}
		return RA_OK;
	};
	std::vector<uint32_t> gf;
	group_boundaries(chains, n_chains, min_score_mode, gf);
	int rc;
	const float *read_dev = ctx->ev_out.as<float>();
	if (read_arena) {
		if ((rc = ctx->a_arena.reserve((size_t)n_read * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->a_arena.p, read_arena, (size_t)n_read * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		read_dev = ctx->a_arena.as<float>();
	}
	const float *ref_dev = ctx->ref.as<float>();
	if (ref_arena) {
		if ((rc = ctx->b_arena.reserve((size_t)n_ref * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->b_arena.p, ref_arena, (size_t)n_ref * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		ref_dev = ctx->b_arena.as<float>();
	}
	if ((rc = ctx->anchors.reserve((size_t)n_anchors * sizeof(ra_anchor_t)))) return rc;
	if ((rc = ctx->chains.reserve((size_t)n_chains * sizeof(ra_chain_t)))) return rc;
	if ((rc = ctx->group_first.reserve(gf.size() * sizeof(uint32_t)))) return rc;
	if ((rc = ctx->score.reserve((size_t)n_chains * sizeof(float)))) return rc;
	if ((rc = ctx->keep.reserve((size_t)n_chains))) return rc;
	const uint32_t n_groups = (uint32_t)gf.size() - 1;
	const float *ms_dev = nullptr;
	if (min_score_mode != 1 && min_score) {
		if ((rc = ctx->min_score.reserve((size_t)n_chains * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->min_score.p, min_score, (size_t)n_chains * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		ms_dev = ctx->min_score.as<float>();
	}
	uint64_t cells[2] = {0, 0};
	const bool want_cells = cells_issued || cells_useful;
	// Sparse border with chains laid out in order: anchors and chain records go up in slices (h2d stream) and every slice is
	// evaluated as soon as it has arrived; anything else takes the one-shot path.
	const bool in_order = laid_in_order && ctx->opt.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_SPARSE;
	const int n_slices = in_order ? (int)std::min<uint64_t>(kMaxSlices, std::max<uint64_t>(1, n_anchors / ctx->chain_slice_min_anchors)) : 1;   // >= 16 MB of anchors per slice by default
	if (n_slices > 1) {
		if ((rc = ctx->counter.reserve(256))) return rc;
		if ((rc = ctx->cells.reserve(64))) return rc;
		if ((rc = ctx->pin_gf.reserve(gf.size() * sizeof(uint32_t)))) return rc;
		memcpy(ctx->pin_gf.p, gf.data(), gf.size() * sizeof(uint32_t));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));              // previous users of the staging buffers are done; the events are resident
		RA_CUDA(cudaMemcpyAsync(ctx->group_first.p, ctx->pin_gf.p, gf.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->h2d));
		uint32_t gcut[kMaxSlices + 1]; gcut[0] = 0;
		{
			uint32_t g = 0;
			for (int k = 1; k <= n_slices; ++k) {
				const uint64_t target = n_anchors * k / n_slices;
				while (g < n_groups && chains[gf[g + 1] - 1].anchor_off + chains[gf[g + 1] - 1].n_anchors <= target) ++g;
				if (k == n_slices) g = n_groups;
				gcut[k] = g;
			}
		}
		for (int k = 0; k < n_slices; ++k) {
			const uint32_t k0 = gf[gcut[k]], k1 = gf[gcut[k + 1]];
			if (k1 > k0) {
				const uint64_t a0 = chains[k0].anchor_off, a1 = chains[k1 - 1].anchor_off + chains[k1 - 1].n_anchors;
				RA_CUDA(cudaMemcpyAsync(ctx->anchors.as<ra_anchor_t>() + a0, anchors + a0, (size_t)(a1 - a0) * sizeof(ra_anchor_t), cudaMemcpyHostToDevice, ctx->h2d));
				RA_CUDA(cudaMemcpyAsync(ctx->chains.as<ra_chain_t>() + k0, chains + k0, (size_t)(k1 - k0) * sizeof(ra_chain_t), cudaMemcpyHostToDevice, ctx->h2d));
			}
			RA_CUDA(cudaEventRecord(ctx->slice_in[k], ctx->h2d));
		}
		tr.mark("chains: copies queued");
		const ra_mapopt_t &o = ctx->opt;
		ra::ChainLaunch L; memset(&L, 0, sizeof(L));
		L.read_arena = read_dev; L.ref_arena = ref_dev; L.anchors = ctx->anchors.as<ra_anchor_t>(); L.chains = ctx->chains.as<ra_chain_t>();
		L.border = o.dtw_border_constraint; L.fill = o.dtw_fill_method; L.fp_contract = o.fp_contract;
		L.band_frac = o.dtw_band_radius_frac; L.match_bonus = o.dtw_match_bonus; L.min_dtw_score = o.dtw_min_score;
		L.min_score_mode = min_score_mode; L.min_score = ms_dev; L.score = ctx->score.as<float>(); L.keep = ctx->keep.as<uint8_t>();
		L.work_counter = ctx->counter.as<uint32_t>(); L.need_cap = ctx->counter.as<uint32_t>() + 4;
		L.error = ctx->counter.as<uint32_t>() + 20; L.n_read = nr; L.n_ref = nf;
		RA_CUDA(cudaMemsetAsync(L.error, 0, sizeof(uint32_t), ctx->stream));
		L.cells = want_cells ? ctx->cells.as<uint64_t>() : nullptr;
		const size_t warps = (size_t)ctx->n_sm * ra::kChainCtasPerSm * 4;
		uint32_t ring_cap = ctx->chain_ring.cap ? (uint32_t)(ctx->chain_ring.cap / (warps * 3 * sizeof(float))) : 0;
		{ uint32_t c = 1; while (c * 2 <= ring_cap) c *= 2; ring_cap = ring_cap ? c : 0; }
		L.ring_scratch = ring_cap ? ctx->chain_ring.as<float>() : nullptr; L.ring_cap = ring_cap;
		RA_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
		for (int k = 0; k < n_slices; ++k) {
			const uint32_t g0 = gcut[k], g1 = gcut[k + 1];
			if ((rc = validate_spans(gf[g0], gf[g1]))) { cudaStreamSynchronize(ctx->h2d); cudaStreamSynchronize(ctx->stream); return rc; }   // overlaps the uploads
			RA_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->slice_in[k], 0));
			if (g1 == g0) continue;
			ra::ChainLaunch Ls = L;
			Ls.group_first = ctx->group_first.as<uint32_t>() + g0; Ls.n_groups = g1 - g0;
			const uint32_t k0 = gf[g0], k1 = gf[g1];
			const uint64_t a0 = chains[k0].anchor_off, a1 = chains[k1 - 1].anchor_off + chains[k1 - 1].n_anchors;
			(void)a0; (void)a1;
			RA_CUDA(ra::launch_validate_chains(Ls, k0, k1, ctx->stream));       // interior anchors of the slice, before anything indexes with them
			ctx->launches += 1;
			if ((rc = sparse_pipeline(ctx, Ls, n_chains, n_anchors, k == 0))) return rc;
		}
		RA_CUDA(cudaEventRecord(ctx->ev[4], ctx->stream));
		ctx->timed_dtw = true;
		tr.mark("chains: kernels queued");
		if (tr.on) { cudaStreamSynchronize(ctx->h2d); tr.mark("chains: h2d done"); cudaStreamSynchronize(ctx->stream); tr.mark("chains: kernels done"); }
		uint32_t need = 0, bad = 0;
		RA_CUDA(cudaMemcpyAsync(&need, L.need_cap, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(&bad, L.error, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		if (L.cells) RA_CUDA(cudaMemcpyAsync(cells, L.cells, sizeof(cells), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(alignment_score, ctx->score.p, (size_t)n_chains * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
		if (keep) RA_CUDA(cudaMemcpyAsync(keep, ctx->keep.p, (size_t)n_chains, cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));
		tr.mark("chains: results back");
		if (bad) return fail(RA_EINVAL, "a chain's anchors are not monotonic (anchors[0] must be the chain end, positions non-increasing) or leave the arenas");
		if (need == 0) {
			if (cells_issued) *cells_issued = cells[0];
			if (cells_useful) *cells_useful = cells[1];
			return RA_OK;
		}
		// a part was wider than the rings at hand: everything is on the device now, take the one-shot path (it grows the ring)
	} else if (ctx->opt.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_GLOBAL && !ctx->force_generic_dtw) {
		// Global border: one region per chain (rmap.cpp:192-237).  The sub-DTWs go through the item kernels (sorted into the
		// register-wavefront classes by size); the pre-check against min_score (:205-209), alignment_score (:306) and the running
		// best_found_alignment of the evaluation loop (:515-524) are replayed here in the reference's order and float arithmetic.
		if ((rc = validate_spans(0, n_chains))) return rc;
		const ra_mapopt_t &o = ctx->opt;
		std::vector<ra_dtw_item_t> items(n_chains);
		for (uint32_t k = 0; k < n_chains; ++k) {
			const ra_chain_t &ch = chains[k];
			const ra_anchor_t &s = anchors[ch.anchor_off + ch.n_anchors - 1], &e = anchors[ch.anchor_off];     // :195-196
			ra_dtw_item_t &it = items[k];
			it.a_off = ch.read_off + s.query_position; it.b_off = ch.ref_off + s.target_position;
			it.a_len = e.query_position - s.query_position + 1; it.b_len = e.target_position - s.target_position + 1;
			it.band_radius = (o.dtw_fill_method == RA_DTW_FILL_METHOD_BANDED) ? std::max(1, (int)((float)it.a_len * o.dtw_band_radius_frac)) : -1;   // :214
			it.exclude_last = 0;
		}
		if ((rc = ctx->items.reserve((size_t)n_chains * sizeof(ra_dtw_item_t)))) return rc;
		if ((rc = ctx->cost.reserve((size_t)n_chains * sizeof(float)))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->items.p, items.data(), (size_t)n_chains * sizeof(ra_dtw_item_t), cudaMemcpyHostToDevice, ctx->stream));
		if ((rc = dtw_dev(ctx, read_dev, ref_dev, ctx->items.as<ra_dtw_item_t>(), items.data(), n_chains, ctx->cost.as<float>()))) return rc;
		std::vector<float> cost(n_chains);
		RA_CUDA(cudaMemcpyAsync(cost.data(), ctx->cost.p, (size_t)n_chains * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
		RA_CUDA(cudaStreamSynchronize(ctx->stream));
		float best = 0.0f;                                                                                     // :515
		for (uint32_t k = 0; k < n_chains; ++k) {
			if (min_score_mode == 1 && (k == 0 || chains[k].group != chains[k - 1].group)) best = 0.0f;
			const float ms = (min_score_mode == 1) ? best : (min_score ? min_score[k] : RA_SCORE_DROPPED);
			const ra_dtw_item_t &it = items[k];
			const uint64_t c = ra_dtw_cells(it.a_len, it.b_len, it.band_radius);
			cells[0] += c;
			float score = RA_SCORE_DROPPED;
			if (!((float)it.a_len * o.dtw_match_bonus < ms)) { score = chain_score(o, it.a_len, cost[k]); cells[1] += c; }   // :205-209, :236, :306
			alignment_score[k] = score;
			const bool kp = score >= o.dtw_min_score;                                                           // :518
			if (keep) keep[k] = kp ? 1 : 0;
			if (kp && score > best) best = score;                                                               // :519-521
		}
		if (cells_issued) *cells_issued = cells[0];
		if (cells_useful) *cells_useful = cells[1];
		return RA_OK;
	} else {
		if ((rc = validate_spans(0, n_chains))) return rc;
		RA_CUDA(cudaMemcpyAsync(ctx->anchors.p, anchors, (size_t)n_anchors * sizeof(ra_anchor_t), cudaMemcpyHostToDevice, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(ctx->chains.p, chains, (size_t)n_chains * sizeof(ra_chain_t), cudaMemcpyHostToDevice, ctx->stream));
		RA_CUDA(cudaMemcpyAsync(ctx->group_first.p, gf.data(), gf.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	}
	if ((rc = chains_dev(ctx, read_dev, ref_dev, ctx->anchors.as<ra_anchor_t>(), ctx->chains.as<ra_chain_t>(), ctx->group_first.as<uint32_t>(),
	                     n_groups, min_score_mode, ms_dev, ctx->score.as<float>(), ctx->keep.as<uint8_t>(), want_cells ? cells : nullptr, n_chains, n_anchors, nr, nf))) return rc;
	RA_CUDA(cudaMemcpyAsync(alignment_score, ctx->score.p, (size_t)n_chains * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	if (keep) RA_CUDA(cudaMemcpyAsync(keep, ctx->keep.p, (size_t)n_chains, cudaMemcpyDeviceToHost, ctx->stream));
	RA_CUDA(cudaStreamSynchronize(ctx->stream));
	if (cells_issued) *cells_issued = cells[0];
	if (cells_useful) *cells_useful = cells[1];
	return RA_OK;
}

int ra_align_chains_batch_dev(ra_ctx *ctx, const float *read_arena_dev, const float *ref_arena_dev,
                              const ra_anchor_t *anchors_dev, uint64_t n_anchors, const ra_chain_t *chains_dev_p, uint32_t n_chains,
                              const uint32_t *group_first_dev, uint32_t n_groups, int min_score_mode, const float *min_score_dev,
                              float *alignment_score_dev, uint8_t *keep_dev, uint64_t *cells_issued, uint64_t *cells_useful)
{
	if (!ctx || (n_groups && (!anchors_dev || !chains_dev_p || !group_first_dev || !alignment_score_dev))) return fail(RA_EINVAL, "NULL argument");
	RA_CUDA(cudaSetDevice(ctx->device));
	const float *read_dev = read_arena_dev ? read_arena_dev : ctx->ev_out.as<float>();
	const float *ref_dev = ref_arena_dev ? ref_arena_dev : ctx->ref.as<float>();
	if (n_groups && (!read_dev || !ref_dev)) return fail(RA_EINVAL, "no read or reference arena on the device");
	uint64_t cells[2] = {0, 0};
	int rc = chains_dev(ctx, read_dev, ref_dev, anchors_dev, chains_dev_p, group_first_dev, n_groups, min_score_mode, min_score_dev,
	                    alignment_score_dev, keep_dev, (cells_issued || cells_useful) ? cells : nullptr, n_chains, n_anchors,
	                    read_arena_dev ? 0 : ctx->resident_events, ref_arena_dev ? 0 : ctx->ref_n);   // caller-owned arenas: sizes unknown here
	if (rc) return rc;
	if (cells_issued) *cells_issued = cells[0];
	if (cells_useful) *cells_useful = cells[1];
	return RA_OK;
}

int ra_align_chains_cigar_batch(ra_ctx *ctx, const float *read_arena, uint64_t n_read, const float *ref_arena, uint64_t n_ref,
                                const ra_anchor_t *anchors, uint64_t n_anchors, const ra_chain_t *chains, uint32_t n_chains,
                                float *alignment_score, float *dtw_cost_out, uint64_t *path_i, uint64_t *path_j, float *path_diff,
                                uint64_t path_cap, uint64_t *path_off)
{
	if (!ctx || !path_off || (n_chains && (!anchors || !chains || !alignment_score || !path_i || !path_j)))
		return fail(RA_EINVAL, "NULL argument");                   // read_arena == NULL: the events resident on the device
	const ra_mapopt_t &o = ctx->opt;
	std::vector<ra_dtw_item_t> items;
	std::vector<uint64_t> first(n_chains + 1);
	for (uint32_t k = 0; k < n_chains; ++k) {
		first[k] = items.size();
		if (chains[k].anchor_off + chains[k].n_anchors > n_anchors) return fail(RA_EINVAL, "chain %u: anchors out of range", k);
		int rc = chain_items(o, anchors + chains[k].anchor_off, chains[k], true, items);
		if (rc) return rc;
	}
	first[n_chains] = items.size();
	uint64_t cap = 0;
	for (const ra_dtw_item_t &it : items) cap += (uint64_t)it.a_len + it.b_len - 1;
	std::vector<float> cost(items.size()), pd(path_diff ? cap : 0);
	std::vector<uint32_t> pi(cap), pj(cap);
	std::vector<uint64_t> poff(items.size() + 1);
	int rc = ra_dtw_tb_batch(ctx, read_arena, n_read, ref_arena, n_ref, items.data(), (uint32_t)items.size(), cost.data(),
	                         pi.data(), pj.data(), path_diff ? pd.data() : nullptr, cap, poff.data());
	if (rc) return rc;
	uint64_t w = 0;
	for (uint32_t k = 0; k < n_chains; ++k) {
		const ra_chain_t &ch = chains[k];
		const ra_anchor_t *anc = anchors + ch.anchor_off;
		path_off[k] = w;
		float dtw_cost = 0.0f; uint32_t n_aligned = 0;
		const uint64_t f = first[k], l = first[k + 1];
		for (uint64_t q = f; q < l; ++q) {
			const uint64_t len = poff[q + 1] - poff[q];
			if (w + len > path_cap) return fail(RA_EINVAL, "path capacity %llu too small", (unsigned long long)path_cap);
			// start anchor of this part: global -> anchors[n-1]; sparse part p -> anchors[parts-p]
			const ra_anchor_t &s = (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_GLOBAL) ? anc[ch.n_anchors - 1]
			                                                                                    : anc[(ch.n_anchors - 1) - (q - f)];
			for (uint64_t t = 0; t < len; ++t) {
				uint64_t pi64 = pi[poff[q] + t], pj64 = pj[poff[q] + t];
				if (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_SPARSE) { pi64 += s.query_position; pj64 += s.target_position; } // :287-288
				path_i[w + t] = pi64; path_j[w + t] = pj64;
				if (path_diff) path_diff[w + t] = pd[poff[q] + t];
			}
			if (o.dtw_border_constraint == RA_DTW_BORDER_CONSTRAINT_GLOBAL && len) {
				// rmap.cpp:230-233: the loop shifts alignment.back() once per element
				path_i[w + len - 1] += len * (uint64_t)s.query_position;
				path_j[w + len - 1] += len * (uint64_t)s.target_position;
			}
			w += len;
			dtw_cost += cost[q];                                                             // :226, :290
			n_aligned += items[q].a_len;                                                     // :236, :292
		}
		if (dtw_cost_out) dtw_cost_out[k] = dtw_cost;
		alignment_score[k] = chain_score(o, n_aligned, dtw_cost);                            // :306
	}
	path_off[n_chains] = w;
	return RA_OK;
}

} // extern "C"
