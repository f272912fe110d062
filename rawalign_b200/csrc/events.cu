// events.cu -- K1: batched event detection, bit-identical to detect_events() of the reference
// (src/revent.c:190-210; arithmetic of :22-32, :34-75, :77-138, :140-188).
//
// What the reference semantics force: prefix_sum / prefix_sum_square are *sequential* fp32
// accumulations (rounding depends on the order), and the peak detector is a sequential two-detector
// state machine.  Everything else (the two t-statistic tracks, the per-event means, the
// normalisation) is position-independent.  So the chunk is processed by five kernels that each use
// the parallelism their stage has:
//
//   k_prefix   one thread per chunk      sequential fp32 scan, float4 in / float4 out       revent.c:22-32
//   k_tstat    one thread per sample     both windows, fully parallel, coalesced            revent.c:34-75
//   k_peaks    one thread per chunk      the two-detector FSM over the t-stat tracks        revent.c:77-138
//   k_scan     one CTA                   CSR offsets from the per-chunk event counts
//   k_finish   one warp per chunk        event means (parallel), fp64 sum/sum2 in array
//                                        order (lane 0), z-normalise (parallel), CSR write  revent.c:140-188
//
// Intermediates (prefix sums, t-stat tracks, peak lists) live in an HBM scratch arena that the
// 126 MB L2 mostly absorbs for sub-batches of a few thousand chunks.
//
// Compiled with -fmad=false: every fused multiply-add below is spelled out and exists only where
// the reference *build* has one (GCC -O3 on an FMA host contracts revent.c:30, :61 twice, :180),
// selected by the CONTRACT template flag; nothing else may be fused.
#include "ra_common.cuh"
#include <float.h>

namespace ra {

// ---------------------------------------------------------------------------------------------------
// k_prefix: ps[0]=pss[0]=0; ps[i+1]=ps[i]+x[i]; pss[i+1]=pss[i]+x[i]*x[i]  (fused when CONTRACT)
// Layout per chunk: ps[0..len] at ps_base + chunk*pstride (pstride % 4 == 0).
// ---------------------------------------------------------------------------------------------------
template <bool CONTRACT>
__global__ void __launch_bounds__(64)
k_prefix(const float *__restrict__ sig, const uint64_t *__restrict__ chunk_off, const uint32_t *__restrict__ chunk_len,
         uint32_t n_chunks, float *__restrict__ ps_base, float *__restrict__ pss_base, uint32_t pstride)
{
	const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= n_chunks) return;
	const uint32_t len = chunk_len[c];
	const float *__restrict__ x = sig + chunk_off[c];
	float *__restrict__ ps = ps_base + (size_t)c * pstride;
	float *__restrict__ pss = pss_base + (size_t)c * pstride;
	float s = 0.0f, q = 0.0f;
	// outputs are written as float4 at positions [4k .. 4k+3]; position 0 holds the leading zero, so
	// the vector written at step k is {P(4k), P(4k+1), P(4k+2), P(4k+3)} with P(0)=0
	float s0 = 0.0f, q0 = 0.0f;      // P(4k) carried from the previous group
	const bool vec = ((reinterpret_cast<uintptr_t>(x) & 15u) == 0);
	uint32_t t = 0;
	if (vec) {
		constexpr int G = 4;         // float4 loads in flight per thread
		for (; t + 4 * G <= len; t += 4 * G) {
			float4 v[G];
#pragma unroll
			for (int g = 0; g < G; ++g) v[g] = __ldcs(reinterpret_cast<const float4 *>(x + t) + g);
#pragma unroll
			for (int g = 0; g < G; ++g) {
				float4 a, b;
				a.x = s0; b.x = q0;
				s = __fadd_rn(s, v[g].x); q = CONTRACT ? __fmaf_rn(v[g].x, v[g].x, q) : __fadd_rn(q, __fmul_rn(v[g].x, v[g].x)); a.y = s; b.y = q;
				s = __fadd_rn(s, v[g].y); q = CONTRACT ? __fmaf_rn(v[g].y, v[g].y, q) : __fadd_rn(q, __fmul_rn(v[g].y, v[g].y)); a.z = s; b.z = q;
				s = __fadd_rn(s, v[g].z); q = CONTRACT ? __fmaf_rn(v[g].z, v[g].z, q) : __fadd_rn(q, __fmul_rn(v[g].z, v[g].z)); a.w = s; b.w = q;
				s = __fadd_rn(s, v[g].w); q = CONTRACT ? __fmaf_rn(v[g].w, v[g].w, q) : __fadd_rn(q, __fmul_rn(v[g].w, v[g].w)); s0 = s; q0 = q;
				*reinterpret_cast<float4 *>(ps + t + 4 * g) = a;
				*reinterpret_cast<float4 *>(pss + t + 4 * g) = b;
			}
		}
	}
	// tail / unaligned path: scalar
	ps[t] = s; pss[t] = q;
	for (; t < len; ++t) {
		const float xv = __ldg(x + t);
		s = __fadd_rn(s, xv);
		q = CONTRACT ? __fmaf_rn(xv, xv, q) : __fadd_rn(q, __fmul_rn(xv, xv));
		ps[t + 1] = s; pss[t + 1] = q;
	}
}

// ---------------------------------------------------------------------------------------------------
// t-statistic of revent.c:50-69 for one window, from prefix values around position p:
//   s_m = ps[p-w], s_0 = ps[p], s_p = ps[p+w]  (same for the squares)
// ---------------------------------------------------------------------------------------------------
template <bool CONTRACT>
__device__ __forceinline__ float tstat_point(float s_m, float s_0, float s_p, float q_m, float q_0, float q_p, float wf)
{
	// at p == w the reference skips the subtraction (revent.c:53-56); ps[0] = pss[0] = +0.0f so
	// x - 0.0f == x bit for bit and the guard is a numerical no-op.
	const float sum1 = __fsub_rn(s_0, s_m);
	const float sumsq1 = __fsub_rn(q_0, q_m);
	const float sum2 = __fsub_rn(s_p, s_0);
	const float sumsq2 = __fsub_rn(q_p, q_0);
	const float mean1 = __fdiv_rn(sum1, wf);
	const float mean2 = __fdiv_rn(sum2, wf);
	const float e1 = __fdiv_rn(sumsq1, wf);
	const float e2 = __fdiv_rn(sumsq2, wf);
	float var;
	if (CONTRACT) {
		const float t1 = __fmaf_rn(-mean1, mean1, e1);           // vfnmadd: e1 - mean1*mean1
		const float t2 = __fadd_rn(t1, e2);
		var = __fmaf_rn(-mean2, mean2, t2);                      // vfnmadd: t2 - mean2*mean2
	} else {
		const float t1 = __fsub_rn(e1, __fmul_rn(mean1, mean1));
		const float t2 = __fadd_rn(t1, e2);
		var = __fsub_rn(t2, __fmul_rn(mean2, mean2));
	}
	var = fmaxf(var, FLT_MIN);                                   // :63
	const float delta_mean = __fsub_rn(mean2, mean1);            // :68
	const float vw = __fdiv_rn(var, wf);
	// :69 -- C fabs()/sqrt() are double; the quotient is rounded to double, then to float
	return __double2float_rn(__ddiv_rn(fabs((double)delta_mean), __dsqrt_rn((double)vw)));
}

// One thread per (chunk, position).  Grid: x = position tiles, y = chunk.
template <bool CONTRACT>
__global__ void __launch_bounds__(256)
k_tstat(const uint32_t *__restrict__ chunk_len, const float *__restrict__ ps_base, const float *__restrict__ pss_base,
        uint32_t pstride, float *__restrict__ t1_base, float *__restrict__ t2_base, uint32_t tstride, uint32_t w1, uint32_t w2)
{
	const uint32_t c = blockIdx.y;
	const uint32_t len = chunk_len[c];
	const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= len) return;
	const float *__restrict__ ps = ps_base + (size_t)c * pstride;
	const float *__restrict__ pss = pss_base + (size_t)c * pstride;
	float v1 = 0.0f, v2 = 0.0f;
	const float s_0 = __ldg(ps + p), q_0 = __ldg(pss + p);
	if (len >= 2u * w1 && w1 >= 2u && p >= w1 && p + w1 <= len)          // :46, :50
		v1 = tstat_point<CONTRACT>(__ldg(ps + p - w1), s_0, __ldg(ps + p + w1), __ldg(pss + p - w1), q_0, __ldg(pss + p + w1), (float)w1);
	if (len >= 2u * w2 && w2 >= 2u && p >= w2 && p + w2 <= len)
		v2 = tstat_point<CONTRACT>(__ldg(ps + p - w2), s_0, __ldg(ps + p + w2), __ldg(pss + p - w2), q_0, __ldg(pss + p + w2), (float)w2);
	t1_base[(size_t)c * tstride + p] = v1;
	t2_base[(size_t)c * tstride + p] = v2;
}

// ---------------------------------------------------------------------------------------------------
// k_peaks: gen_peaks, revent.c:77-138
// ---------------------------------------------------------------------------------------------------
struct Detector {
	uint32_t masked_to;
	int peak_pos;
	float peak_value;
	int valid_peak;
};

// One detector step, revent.c:89-133.  SHORT selects the masking side effect of :113-120.
template <bool SHORT>
__device__ __forceinline__ void detector_step(Detector &d, Detector &long_d, uint32_t i, float v, float threshold,
                                              uint32_t window_length, float peak_height, uint32_t *__restrict__ peaks, uint32_t &n_peaks)
{
	if (d.masked_to >= i) return;                                 // :91 (unsigned compare: i = 0 is always skipped)
	if (d.peak_pos == -1) {                                       // :94
		if (v < d.peak_value) d.peak_value = v;
		else if (__fsub_rn(v, d.peak_value) > peak_height) { d.peak_value = v; d.peak_pos = (int)i; }
	} else {
		if (v > d.peak_value) { d.peak_value = v; d.peak_pos = (int)i; }   // :107-111
		if (SHORT && d.peak_value > threshold) {                  // :113-120
			long_d.masked_to = (uint32_t)d.peak_pos + window_length;
			long_d.peak_pos = -1; long_d.peak_value = FLT_MAX; long_d.valid_peak = 0;
		}
		if (__fsub_rn(d.peak_value, v) > peak_height && d.peak_value > threshold) d.valid_peak = 1; // :122
		if (d.valid_peak && (i - (uint32_t)d.peak_pos) > window_length / 2) {                         // :126
			peaks[n_peaks++] = (uint32_t)d.peak_pos;              // :128
			d.peak_pos = -1; d.peak_value = v; d.valid_peak = 0;
		}
	}
}

__global__ void __launch_bounds__(32)
k_peaks(const uint32_t *__restrict__ chunk_len, uint32_t n_chunks, const float *__restrict__ t1_base,
        const float *__restrict__ t2_base, uint32_t tstride, EvParams prm, uint32_t *__restrict__ peaks_base,
        uint32_t peak_stride, uint32_t *__restrict__ n_peaks_out)
{
	const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= n_chunks) return;
	const uint32_t len = chunk_len[c];
	const float *__restrict__ t1 = t1_base + (size_t)c * tstride;
	const float *__restrict__ t2 = t2_base + (size_t)c * tstride;
	uint32_t *__restrict__ peaks = peaks_base + (size_t)c * peak_stride;
	Detector sd = { 0u, -1, FLT_MAX, 0 };                          // :198-202
	Detector ld = { 0u, -1, FLT_MAX, 0 };
	uint32_t n_peaks = 0;
	uint32_t i = 0;
	// tstride % 4 == 0 and the tracks are 16-byte aligned: stream them as float4, two groups ahead
	const uint32_t len4 = len & ~3u;
	if (len4 >= 8) {
		float4 a = __ldcs(reinterpret_cast<const float4 *>(t1)), b = __ldcs(reinterpret_cast<const float4 *>(t2));
		float4 a1 = __ldcs(reinterpret_cast<const float4 *>(t1 + 4)), b1 = __ldcs(reinterpret_cast<const float4 *>(t2 + 4));
		for (; i < len4; i += 4) {
			float4 a2 = a1, b2 = b1;
			if (i + 8 < len4) { a2 = __ldcs(reinterpret_cast<const float4 *>(t1 + i + 8)); b2 = __ldcs(reinterpret_cast<const float4 *>(t2 + i + 8)); }
			detector_step<true>(sd, ld, i, a.x, prm.thr1, prm.w1, prm.peak_height, peaks, n_peaks);
			detector_step<false>(ld, ld, i, b.x, prm.thr2, prm.w2, prm.peak_height, peaks, n_peaks);
			detector_step<true>(sd, ld, i + 1, a.y, prm.thr1, prm.w1, prm.peak_height, peaks, n_peaks);
			detector_step<false>(ld, ld, i + 1, b.y, prm.thr2, prm.w2, prm.peak_height, peaks, n_peaks);
			detector_step<true>(sd, ld, i + 2, a.z, prm.thr1, prm.w1, prm.peak_height, peaks, n_peaks);
			detector_step<false>(ld, ld, i + 2, b.z, prm.thr2, prm.w2, prm.peak_height, peaks, n_peaks);
			detector_step<true>(sd, ld, i + 3, a.w, prm.thr1, prm.w1, prm.peak_height, peaks, n_peaks);
			detector_step<false>(ld, ld, i + 3, b.w, prm.thr2, prm.w2, prm.peak_height, peaks, n_peaks);
			a = a1; b = b1; a1 = a2; b1 = b2;
		}
	}
	for (; i < len; ++i) {
		detector_step<true>(sd, ld, i, __ldg(t1 + i), prm.thr1, prm.w1, prm.peak_height, peaks, n_peaks);
		detector_step<false>(ld, ld, i, __ldg(t2 + i), prm.thr2, prm.w2, prm.peak_height, peaks, n_peaks);
	}
	n_peaks_out[c] = n_peaks;      // == number of events (see k_finish)
}

// ---------------------------------------------------------------------------------------------------
// Exclusive prefix sum of per-chunk event counts -> CSR offsets (single CTA; n is the number of
// chunks of one sub-batch).  With chained != 0 the scan continues from out[0] (the total the
// previous sub-batch left there).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_exclusive_scan_u32(const uint32_t *__restrict__ in, uint64_t *__restrict__ out, uint32_t n, int chained)
{
	__shared__ uint64_t warp_tot[32];
	__shared__ uint64_t carry_s, block_tot;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	if (threadIdx.x == 0) carry_s = chained ? out[0] : 0;
	__syncthreads();
	for (uint32_t base = 0; base < n; base += blockDim.x) {
		const uint32_t idx = base + threadIdx.x;
		uint64_t v = idx < n ? in[idx] : 0, incl = v;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { uint64_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
		if (lane == 31) warp_tot[wid] = incl;
		__syncthreads();
		if (wid == 0) {
			const uint64_t w = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;   // any CTA size that is a multiple of 32
			uint64_t wi = w;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { uint64_t y = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += y; }
			warp_tot[lane] = wi - w;                               // exclusive over warps
			if (lane == 31) block_tot = wi;
		}
		__syncthreads();
		if (idx < n) out[idx] = carry_s + warp_tot[wid] + incl - v;
		__syncthreads();
		if (threadIdx.x == 0) carry_s += block_tot;
		__syncthreads();
	}
	if (threadIdx.x == 0) out[n] = carry_s;
}

// ---------------------------------------------------------------------------------------------------
// k_finish: gen_events, revent.c:140-188.  One warp per chunk.
//
// gen_events consumes peaks[0 .. n_ev-2] in emission order and closes the last event at s_len
// (:160-172): the final peak of a chunk never bounds an event.  n_ev == n_peaks because the filter
// of :145-147 (0 < peaks[i] < s_len) always holds: position 0 is never examined (masked_to(0) >= 0,
// :91) and positions are < s_len by construction.
// ---------------------------------------------------------------------------------------------------
template <bool CONTRACT>
__global__ void __launch_bounds__(128)
k_finish(const uint32_t *__restrict__ chunk_len, uint32_t n_chunks, const float *__restrict__ ps_base, uint32_t pstride,
         const uint32_t *__restrict__ peaks_base, uint32_t peak_stride, const uint32_t *__restrict__ n_peaks_in,
         const uint64_t *__restrict__ ev_off, float *__restrict__ ev_out, uint64_t ev_cap, float *__restrict__ raw_base,
         uint32_t *__restrict__ ev_pos)
{
	const uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (c >= n_chunks) return;
	const uint32_t n_ev = n_peaks_in[c];
	if (n_ev == 0) return;                                         // detect_events returns NULL, *n untouched (:205-206)
	const uint32_t len = chunk_len[c];
	const float *__restrict__ ps = ps_base + (size_t)c * pstride;
	const uint32_t *__restrict__ peaks = peaks_base + (size_t)c * peak_stride;
	float *__restrict__ raw = raw_base + (size_t)c * peak_stride;  // un-normalised means (scratch, L2-resident)
	// event k: (ps[end_k] - ps[start_k]) / ((float)end_k - (float)start_k), end_k = peaks[k] (k < n_ev-1) or len,
	// start_k = peaks[k-1] or 0                                   :160-172
	for (uint32_t k = lane; k < n_ev; k += 32) {
		const uint32_t e_pos = (k + 1 < n_ev) ? peaks[k] : len;
		const uint32_t s_pos = (k > 0) ? peaks[k - 1] : 0u;
		const float l_ps = (k > 0) ? __ldg(ps + s_pos) : 0.0f;
		const float l_peak = (k > 0) ? __uint2float_rn(s_pos) : 0.0f;
		raw[k] = __fdiv_rn(__fsub_rn(__ldg(ps + e_pos), l_ps), __fsub_rn(__uint2float_rn(e_pos), l_peak));
	}
	__syncwarp();
	// sum / sum2 in array order, fp64 (:163-164, :173-174); lanes fetch 32 values at a time, lane 0 accumulates
	double sum = 0.0, sum2 = 0.0;
	for (uint32_t base = 0; base < n_ev; base += 32) {
		const float mine = (base + lane < n_ev) ? raw[base + lane] : 0.0f;
		const uint32_t cnt = min(32u, n_ev - base);
		for (uint32_t t = 0; t < cnt; ++t) {
			const float e = __shfl_sync(0xffffffffu, mine, (int)t);
			sum = __dadd_rn(sum, (double)e);
			sum2 = __dadd_rn(sum2, (double)__fmul_rn(e, e));       // float product, then widened (:164)
		}
	}
	const double mean = __ddiv_rn(sum, (double)n_ev);              // :179
	const double q = __ddiv_rn(sum2, (double)n_ev);
	const double var = CONTRACT ? __fma_rn(-mean, mean, q) : __dsub_rn(q, __dmul_rn(mean, mean));   // :180
	const double sd = __dsqrt_rn(var);
	const uint64_t off = ev_off[c];
	for (uint32_t k = lane; k < n_ev; k += 32)
		if (off + k < ev_cap) {
			ev_out[off + k] = __double2float_rn(__ddiv_rn(__dsub_rn((double)raw[k], mean), sd));        // :183
			if (ev_pos) ev_pos[off + k] = (k > 0) ? peaks[k - 1] : 0u;   // first sample of the event (not part of the reference's output)
		}
}

// ---------------------------------------------------------------------------------------------------
bool events_windows_supported(uint32_t w1, uint32_t w2) { return w1 >= 1 && w2 >= 1 && w1 < 4096 && w2 < 4096; }

void events_scratch_layout(uint32_t max_len, const EvParams &p, EvScratch *s)
{
	s->pstride = ((max_len + 1 + 3) & ~3u) + 4;
	s->tstride = ((max_len + 3) & ~3u) + 4;
	// a detector emits at i >= peak_pos + w/2 + 1 and can only start its next peak at a later sample,
	// so peaks of one detector are >= w/2 + 2 apart
	s->peak_stride = max_len / (p.w1 / 2 + 2) + max_len / (p.w2 / 2 + 2) + 4;
	s->floats_per_chunk = 2 * (size_t)s->pstride + 2 * (size_t)s->tstride + 2 * (size_t)s->peak_stride;
}

cudaError_t launch_events(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                          uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                          uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, cudaStream_t s,
                          cudaEvent_t after_detect)
{
	if (n_chunks == 0) return cudaSuccess;
	EvScratch L; events_scratch_layout(max_len, p, &L);
	float *ps = scratch;
	float *pss = ps + (size_t)n_chunks * L.pstride;
	float *t1 = pss + (size_t)n_chunks * L.pstride;
	float *t2 = t1 + (size_t)n_chunks * L.tstride;
	uint32_t *peaks = reinterpret_cast<uint32_t *>(t2 + (size_t)n_chunks * L.tstride);
	float *raw = reinterpret_cast<float *>(peaks + (size_t)n_chunks * L.peak_stride);
	const dim3 tgrid((max_len + 255) / 256, n_chunks);
	if (contract) {
		k_prefix<true><<<(n_chunks + 63) / 64, 64, 0, s>>>(sig, chunk_off, chunk_len, n_chunks, ps, pss, L.pstride);
		k_tstat<true><<<tgrid, 256, 0, s>>>(chunk_len, ps, pss, L.pstride, t1, t2, L.tstride, p.w1, p.w2);
	} else {
		k_prefix<false><<<(n_chunks + 63) / 64, 64, 0, s>>>(sig, chunk_off, chunk_len, n_chunks, ps, pss, L.pstride);
		k_tstat<false><<<tgrid, 256, 0, s>>>(chunk_len, ps, pss, L.pstride, t1, t2, L.tstride, p.w1, p.w2);
	}
	k_peaks<<<(n_chunks + 31) / 32, 32, 0, s>>>(chunk_len, n_chunks, t1, t2, L.tstride, p, peaks, L.peak_stride, n_ev);
	if (after_detect) cudaEventRecord(after_detect, s);
	k_exclusive_scan_u32<<<1, 1024, 0, s>>>(n_ev, ev_off, n_chunks, chained);
	const unsigned fblocks = (unsigned)(((uint64_t)n_chunks * 32 + 127) / 128);
	if (contract) k_finish<true><<<fblocks, 128, 0, s>>>(chunk_len, n_chunks, ps, L.pstride, peaks, L.peak_stride, n_ev, ev_off, ev_out, ev_cap, raw, ev_pos);
	else k_finish<false><<<fblocks, 128, 0, s>>>(chunk_len, n_chunks, ps, L.pstride, peaks, L.peak_stride, n_ev, ev_off, ev_out, ev_cap, raw, ev_pos);
	return cudaGetLastError();
}

cudaError_t launch_exclusive_scan_u32_chained(const uint32_t *in, uint64_t *out, uint32_t n, int chained, cudaStream_t s)
{
	k_exclusive_scan_u32<<<1, 1024, 0, s>>>(in, out, n, chained);
	return cudaGetLastError();
}

cudaError_t launch_exclusive_scan_u32(const uint32_t *in, uint64_t *out, uint32_t n, cudaStream_t s)
{
	k_exclusive_scan_u32<<<1, 1024, 0, s>>>(in, out, n, 0);
	return cudaGetLastError();
}

} // namespace ra
