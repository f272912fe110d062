// noalloc.cu -- ra_detect_events_noalloc: event detection (K1, the v4 pipeline of events_pipe.cu) for callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The pipeline's scratch (prefix sums, the two t-statistic tracks, peak lists, per-chunk counts) lives in ONE file-scope device
// array of kNoallocScratchBytes that every GPU which loads the library holds until exit.  The entry point calls no cudaMalloc,
// cudaFree or allocating library routine; it enqueues the five kernels on the caller's stream and returns without synchronising.
// Every call shares the array: calls that overlap on the GPU (two streams, two host threads, two contexts on one device) race.
// The kernels write every scratch word before they read it, so stale contents of the previous call are harmless; the per-chunk
// event counts are cleared anyway.
#include "ra_common.cuh"
#include <mutex>

extern "C" int ra__device(const ra_ctx *ctx);
extern "C" int ra__fail(int code, const char *fmt, ...);
extern "C" void ra__add_launches(ra_ctx *ctx, uint64_t n);

namespace {

constexpr size_t kNoallocScratchBytes = 8u << 20;                  // 8 MiB per GPU: e.g. 100 chunks of 4096 samples with the default windows (3, 6)
constexpr size_t kNoallocScratchFloats = kNoallocScratchBytes / sizeof(float);
static __device__ __align__(256) float g_noalloc_scratch[kNoallocScratchFloats];

// In host code the array's name is a host-side stand-in, not a device address: the address is asked for once per device, with
// that device current, and only that pointer is handed to the kernels.
constexpr int kMaxDevices = 64;
float *g_scratch_dev[kMaxDevices] = {};
std::mutex g_scratch_mutex;

int scratch_for_current_device(int dev, float **out)
{
	if (dev < 0 || dev >= kMaxDevices) return ra__fail(RA_EUNSUP, "device ordinal %d (supported: 0..%d)", dev, kMaxDevices - 1);
	std::lock_guard<std::mutex> lock(g_scratch_mutex);
	if (!g_scratch_dev[dev]) {
		float *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_scratch);
		if (e != cudaSuccess || !p) return ra__fail(RA_ECUDA, "cudaGetSymbolAddress(noalloc scratch) on device %d -> %s", dev, cudaGetErrorString(e));
		g_scratch_dev[dev] = p;
	}
	*out = g_scratch_dev[dev];
	return RA_OK;
}

} // namespace

extern "C" int ra_detect_events_noalloc(ra_ctx *ctx, const float *sig_dev, const uint64_t *chunk_off_dev, const uint32_t *chunk_len_dev,
                                        uint32_t n_chunks, uint32_t max_chunk_len, float *ev_out_dev, uint64_t ev_cap, uint64_t *ev_off_dev,
                                        void *stream)
{
	if (!ctx || !ev_off_dev || (n_chunks && (!sig_dev || !chunk_off_dev || !chunk_len_dev || !ev_out_dev))) return ra__fail(RA_EINVAL, "NULL argument");
	ra_mapopt_t o;
	int rc = ra_get_opt(ctx, &o);
	if (rc) return rc;
	if (!ra::events_pipe_supported(o.window_length1, o.window_length2))
		return ra__fail(RA_EUNSUP, "event detector windows (%u,%u) are not supported by the no-allocation entry point", o.window_length1, o.window_length2);
	const int dev = ra__device(ctx);
	cudaError_t e = cudaSetDevice(dev);
	if (e != cudaSuccess) return ra__fail(RA_ECUDA, "cudaSetDevice(%d) -> %s", dev, cudaGetErrorString(e));
	cudaStream_t s = static_cast<cudaStream_t>(stream);
	if (n_chunks == 0) {
		e = cudaMemsetAsync(ev_off_dev, 0, sizeof(uint64_t), s);
		return e == cudaSuccess ? RA_OK : ra__fail(RA_ECUDA, "cudaMemsetAsync -> %s", cudaGetErrorString(e));
	}
	const ra::EvParams p = { o.window_length1, o.window_length2, o.threshold1, o.threshold2, o.peak_height };
	uint32_t stride = 0, pstride = 0; size_t fpc = 0;
	ra::events_pipe_layout(max_chunk_len, p, &stride, &pstride, &fpc);
	// layout: [pipeline scratch of n_chunks chunks | per-chunk event counts]
	const size_t need = (size_t)n_chunks * fpc + (((size_t)n_chunks + 31u) & ~(size_t)31u);
	if (need > kNoallocScratchFloats)
		return ra__fail(RA_EINVAL, "%u chunks of up to %u samples need %zu bytes of scratch; the no-allocation entry point holds %zu bytes (%zu chunks of this length)",
		                n_chunks, max_chunk_len, need * sizeof(float), kNoallocScratchBytes, (kNoallocScratchFloats - 32) / (fpc + 1));
	float *scratch = nullptr;
	if ((rc = scratch_for_current_device(dev, &scratch))) return rc;
	uint32_t *n_ev = reinterpret_cast<uint32_t *>(scratch + (size_t)n_chunks * fpc);
	if ((e = cudaMemsetAsync(n_ev, 0, (size_t)n_chunks * sizeof(uint32_t), s)) != cudaSuccess) return ra__fail(RA_ECUDA, "cudaMemsetAsync -> %s", cudaGetErrorString(e));
	e = ra::launch_events_pipe(sig_dev, chunk_off_dev, chunk_len_dev, n_chunks, max_chunk_len, p, (int)o.fp_contract, scratch, n_ev, ev_off_dev, 0,
	                           ev_out_dev, ev_cap, nullptr, s, nullptr);
	if (e != cudaSuccess) return ra__fail(RA_ECUDA, "event pipeline launch -> %s", cudaGetErrorString(e));
	ra__add_launches(ctx, ra::kPipeEventKernels);
	return RA_OK;
}
