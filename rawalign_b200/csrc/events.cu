// events.cu -- K1: batched event detection, bit-identical to detect_events() of the reference
// (src/revent.c:190-210; arithmetic of :22-32, :34-75, :77-138, :140-188).
//
// Parallelism: the reference's prefix sums are *sequential* fp32 accumulations and its peak
// detector is a sequential two-detector state machine, so bit-exactness forces one sequential
// scan per chunk.  One thread owns one chunk and fuses all four reference passes into a single
// sweep: it keeps the last 2*W2+1 prefix values in a register ring (the t-statistic at position
// p needs prefix[p-w], prefix[p], prefix[p+w] only), steps both detectors W2 samples behind the
// scan head, and turns every emitted peak into an event mean on the fly.  Parallelism comes from
// the batch: >= 1e5 chunks in flight.  Each thread streams its own chunk with 16-byte loads
// (float4); the 32 B sector a lane touches is reused from L1 on its next load.
//
// Compiled with -fmad=false: every fused multiply-add below is spelled out and exists only where
// the reference *build* has one (GCC -O3 on an FMA host contracts revent.c:30, :61 twice, :180),
// selected by the CONTRACT template flag; nothing else may be fused.
#include "ra_common.cuh"
#include <float.h>

namespace ra {

template <int W2> struct RingSize { static constexpr int value = (2 * W2 + 1 <= 16) ? 16 : ((2 * W2 + 1 <= 32) ? 32 : 64); };

// t-statistic of revent.c:50-69 for one window, from prefix values around position p.
//   s_m = ps[p-w], s_0 = ps[p], s_p = ps[p+w]  (same for the squares)
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

struct Detector {
	uint32_t masked_to;
	int peak_pos;
	float peak_value;
	int valid_peak;
	float peak_ps;     // prefix_sum[peak_pos], captured when peak_pos is assigned (gen_events needs it later)
};

struct EventAcc {
	float l_ps, l_peak;         // revent.c:155
	double sum, sum2;           // :143
	uint32_t n_peaks;
	uint32_t pend_pos; float pend_ps;   // the most recent peak; becomes an event only when another peak follows
	float *out;                 // raw (un-normalised) events of this chunk
};

// gen_events consumes peaks[0 .. n_ev-2] in emission order and closes the last event at s_len
// (revent.c:160-172): the final peak of a chunk never bounds an event.  n_ev == n_peaks because the
// filter of :145-147 (0 < peaks[i] < s_len) always holds: position 0 is never examined
// (masked_to(0) >= 0, :91) and positions are < s_len by construction.
__device__ __forceinline__ void emit_peak(EventAcc &E, uint32_t pos, float ps_at)
{
	if (E.n_peaks > 0) {
		const float e = __fdiv_rn(__fsub_rn(E.pend_ps, E.l_ps), __fsub_rn(__uint2float_rn(E.pend_pos), E.l_peak)); // :162
		E.out[E.n_peaks - 1] = e;
		E.sum = __dadd_rn(E.sum, (double)e);                      // :163
		E.sum2 = __dadd_rn(E.sum2, (double)__fmul_rn(e, e));      // :164 (float product, then widened)
		E.l_ps = E.pend_ps;                                       // :165
		E.l_peak = __uint2float_rn(E.pend_pos);                   // :166
	}
	E.pend_pos = pos; E.pend_ps = ps_at;
	E.n_peaks++;
}

// One detector step, revent.c:89-133.  SHORT selects the masking side effect of :113-120.
template <bool SHORT>
__device__ __forceinline__ void detector_step(Detector &d, Detector &long_d, uint32_t i, float v, float ps_i,
                                              float threshold, uint32_t window_length, float peak_height, EventAcc &E)
{
	if (d.masked_to >= i) return;                                 // :91
	if (d.peak_pos == -1) {                                       // :94
		if (v < d.peak_value) d.peak_value = v;
		else if (__fsub_rn(v, d.peak_value) > peak_height) { d.peak_value = v; d.peak_pos = (int)i; d.peak_ps = ps_i; }
	} else {
		if (v > d.peak_value) { d.peak_value = v; d.peak_pos = (int)i; d.peak_ps = ps_i; }   // :107-111
		if (SHORT && d.peak_value > threshold) {                  // :113-120
			long_d.masked_to = (uint32_t)d.peak_pos + window_length;
			long_d.peak_pos = -1; long_d.peak_value = FLT_MAX; long_d.valid_peak = 0;
		}
		if (__fsub_rn(d.peak_value, v) > peak_height && d.peak_value > threshold) d.valid_peak = 1; // :122
		if (d.valid_peak && (i - (uint32_t)d.peak_pos) > window_length / 2) {                         // :126
			emit_peak(E, (uint32_t)d.peak_pos, d.peak_ps);
			d.peak_pos = -1; d.peak_value = v; d.valid_peak = 0;
		}
	}
}

template <int W1, int W2, bool CONTRACT>
__global__ void __launch_bounds__(128)
k_events_scan(const float *__restrict__ sig, const uint64_t *__restrict__ chunk_off,
              const uint32_t *__restrict__ chunk_len, uint32_t n_chunks, EvParams prm,
              float *__restrict__ raw, uint32_t raw_stride, uint32_t *__restrict__ n_ev_out,
              EvChunkStat *__restrict__ stat_out)
{
	constexpr int RING = RingSize<W2>::value;
	constexpr int MASK = RING - 1;
	const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= n_chunks) return;
	const uint32_t s_len = chunk_len[c];
	const float *__restrict__ x = sig + chunk_off[c];
	const bool vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15u) == 0);

	float rs[RING], rq[RING];                                    // prefix_sum / prefix_sum_square ring, index = position & MASK
#pragma unroll
	for (int k = 0; k < RING; ++k) { rs[k] = 0.0f; rq[k] = 0.0f; }
	float ps = 0.0f, pss = 0.0f;                                 // :26-27

	Detector sd = { 0u, -1, FLT_MAX, 0, 0.0f };                   // :198-202
	Detector ld = { 0u, -1, FLT_MAX, 0, 0.0f };
	EventAcc E;
	E.l_ps = 0.0f; E.l_peak = 0.0f; E.sum = 0.0; E.sum2 = 0.0; E.n_peaks = 0; E.pend_pos = 0; E.pend_ps = 0.0f;
	E.out = raw + (size_t)c * raw_stride;

	const bool t1_on = (s_len >= 2u * W1) && (W1 >= 2);           // :46
	const bool t2_on = (s_len >= 2u * W2) && (W2 >= 2);
	const float w1f = (float)W1, w2f = (float)W2;

	// iteration t: (A) extend the prefix sums to position t+1, (B) step the detectors at p = t+1-W2
	const uint32_t t_end = s_len + (uint32_t)W2 - 1u;             // last t has p = s_len-1
	for (uint32_t t0 = 0; t0 < t_end; t0 += RING) {
		float xin[RING];
		if (vec_ok && t0 + RING <= s_len) {
#pragma unroll
			for (int k = 0; k < RING / 4; ++k) {
				const float4 v = __ldg(reinterpret_cast<const float4 *>(x + t0) + k);
				xin[4 * k] = v.x; xin[4 * k + 1] = v.y; xin[4 * k + 2] = v.z; xin[4 * k + 3] = v.w;
			}
		} else {
#pragma unroll
			for (int k = 0; k < RING; ++k) xin[k] = (t0 + k < s_len) ? __ldg(x + t0 + k) : 0.0f;
		}
#pragma unroll
		for (int u = 0; u < RING; ++u) {
			const uint32_t t = t0 + u;                             // t & MASK == u  (t0 is a multiple of RING)
			if (t >= t_end) break;
			if (t < s_len) {                                       // (A) revent.c:28-31
				const float xv = xin[u];
				ps = __fadd_rn(ps, xv);
				pss = CONTRACT ? __fmaf_rn(xv, xv, pss) : __fadd_rn(pss, __fmul_rn(xv, xv));
				rs[(u + 1) & MASK] = ps; rq[(u + 1) & MASK] = pss;
			}
			if (t + 1u >= (uint32_t)W2) {                          // (B)
				const uint32_t p = t + 1u - (uint32_t)W2;          // position & MASK == (u + 1 - W2) & MASK
				constexpr int dummy = 0; (void)dummy;
				const int ip = (u + 1 - W2) & MASK;
				const float s_0 = rs[ip], q_0 = rq[ip];
				float v1 = 0.0f, v2 = 0.0f;
				if (t1_on && p >= (uint32_t)W1 && p + (uint32_t)W1 <= s_len)
					v1 = tstat_point<CONTRACT>(rs[(ip - W1) & MASK], s_0, rs[(ip + W1) & MASK],
					                           rq[(ip - W1) & MASK], q_0, rq[(ip + W1) & MASK], w1f);
				if (t2_on && p >= (uint32_t)W2 && p + (uint32_t)W2 <= s_len)
					v2 = tstat_point<CONTRACT>(rs[(ip - W2) & MASK], s_0, rs[(ip + W2) & MASK],
					                           rq[(ip - W2) & MASK], q_0, rq[(ip + W2) & MASK], w2f);
				detector_step<true>(sd, ld, p, v1, s_0, prm.thr1, (uint32_t)W1, prm.peak_height, E);
				detector_step<false>(ld, ld, p, v2, s_0, prm.thr2, (uint32_t)W2, prm.peak_height, E);
			}
		}
	}

	uint32_t n_ev = 0;
	double mean = 0.0, stdv = 1.0;
	if (E.n_peaks > 0) {                                          // :206
		n_ev = E.n_peaks;
		const float e = __fdiv_rn(__fsub_rn(ps, E.l_ps), __fsub_rn(__uint2float_rn(s_len), E.l_peak)); // :172 (ps == prefix_sum[s_len])
		E.out[n_ev - 1] = e;
		E.sum = __dadd_rn(E.sum, (double)e);
		E.sum2 = __dadd_rn(E.sum2, (double)__fmul_rn(e, e));
		mean = __ddiv_rn(E.sum, (double)n_ev);                     // :179
		const double q = __ddiv_rn(E.sum2, (double)n_ev);
		const double var = CONTRACT ? __fma_rn(-mean, mean, q) : __dsub_rn(q, __dmul_rn(mean, mean)); // :180
		stdv = __dsqrt_rn(var);
	}
	n_ev_out[c] = n_ev;
	stat_out[c].mean = mean; stat_out[c].stdv = stdv;
}

// Exclusive prefix sum of per-chunk event counts -> CSR offsets (single CTA; n is the number of
// chunks of one batch, the scan is a few microseconds of plumbing next to the detector).
__global__ void __launch_bounds__(1024) k_exclusive_scan_u32(const uint32_t *__restrict__ in, uint64_t *__restrict__ out, uint32_t n)
{
	__shared__ uint64_t warp_tot[32];
	__shared__ uint64_t carry_s, block_tot;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	if (threadIdx.x == 0) carry_s = 0;
	__syncthreads();
	for (uint32_t base = 0; base < n; base += blockDim.x) {
		const uint32_t idx = base + threadIdx.x;
		uint64_t v = idx < n ? in[idx] : 0, incl = v;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { uint64_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
		if (lane == 31) warp_tot[wid] = incl;
		__syncthreads();
		if (wid == 0) {
			const uint64_t w = warp_tot[lane];
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

// (ev - mean) / std_dev of revent.c:182-184, written to the CSR output; one warp per chunk.
__global__ void __launch_bounds__(256)
k_events_normalise(const float *__restrict__ raw, uint32_t raw_stride, const uint32_t *__restrict__ n_ev,
                   const EvChunkStat *__restrict__ stat, const uint64_t *__restrict__ ev_off, uint32_t n_chunks,
                   float *__restrict__ ev_out, uint64_t ev_cap)
{
	const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (warp >= n_chunks) return;
	const uint32_t n = n_ev[warp];
	if (n == 0) return;
	const double mean = stat[warp].mean, sd = stat[warp].stdv;
	const float *__restrict__ r = raw + (size_t)warp * raw_stride;
	const uint64_t off = ev_off[warp];
	for (uint32_t k = lane; k < n; k += 32)
		if (off + k < ev_cap)
			ev_out[off + k] = __double2float_rn(__ddiv_rn(__dsub_rn((double)r[k], mean), sd));
}

bool events_windows_supported(uint32_t w1, uint32_t w2) { return (w1 == 3 && w2 == 6) || (w1 == 7 && w2 == 14); }

cudaError_t launch_events_scan(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len,
                               uint32_t n_chunks, const EvParams &p, int contract, float *raw,
                               uint32_t raw_stride, uint32_t *n_ev, EvChunkStat *stat, cudaStream_t s)
{
	if (n_chunks == 0) return cudaSuccess;
	const int threads = 128;
	const unsigned blocks = (n_chunks + threads - 1) / threads;
#define RA_EV_LAUNCH(W1, W2)                                                                                       \
	do {                                                                                                           \
		if (contract) k_events_scan<W1, W2, true><<<blocks, threads, 0, s>>>(sig, chunk_off, chunk_len, n_chunks, p, raw, raw_stride, n_ev, stat); \
		else k_events_scan<W1, W2, false><<<blocks, threads, 0, s>>>(sig, chunk_off, chunk_len, n_chunks, p, raw, raw_stride, n_ev, stat);         \
	} while (0)
	if (p.w1 == 3 && p.w2 == 6) RA_EV_LAUNCH(3, 6);
	else if (p.w1 == 7 && p.w2 == 14) RA_EV_LAUNCH(7, 14);
	else return cudaErrorInvalidValue;
#undef RA_EV_LAUNCH
	return cudaGetLastError();
}

cudaError_t launch_exclusive_scan_u32(const uint32_t *in, uint64_t *out, uint32_t n, cudaStream_t s)
{
	k_exclusive_scan_u32<<<1, 1024, 0, s>>>(in, out, n);
	return cudaGetLastError();
}

cudaError_t launch_events_normalise(const float *raw, uint32_t raw_stride, const uint32_t *n_ev,
                                    const EvChunkStat *stat, const uint64_t *ev_off, uint32_t n_chunks,
                                    float *ev_out, uint64_t ev_cap, cudaStream_t s)
{
	if (n_chunks == 0) return cudaSuccess;
	const int threads = 256;
	const unsigned blocks = (unsigned)(((uint64_t)n_chunks * 32 + threads - 1) / threads);
	k_events_normalise<<<blocks, threads, 0, s>>>(raw, raw_stride, n_ev, stat, ev_off, n_chunks, ev_out, ev_cap);
	return cudaGetLastError();
}

} // namespace ra
