// events_fused.cu -- K1 v3: detect_events() (src/revent.c:190-210) for a batch of chunks in ONE pass over
// the signal.  Bit-identical to the reference build (same sequential fp32 prefix sums, same fused
// multiply-adds where GCC contracts, same fp64 sqrt/div rounding, same detector state machine).
//
// What is sequential per chunk (the fp32 prefix sums, revent.c:22-32, and the two-detector peak state
// machine, :77-138) is run by ONE lane per chunk; what is position-independent (the two t-statistic
// tracks, :34-75) is run by all 32 lanes of the warp over a 32-sample tile.  A warp owns k consecutive
// chunks (k = 1..16, chosen by the launcher so that small batches still fill the machine) and walks
// them tile by tile:
//
//   load     32 samples of each of the k chunks, one coalesced 128-byte row per chunk; the next tile is
//            already in registers while this one is processed (the only HBM round trip on the path)
//   prefix   lane c < k: ps/pss of chunk c for the 32 new samples into 64-entry shared-memory rings
//   t-stat   all lanes, chunk by chunk: both windows for 32 positions (lagging the prefix by w_max);
//            division by the window is a 3-instruction exact sequence (checked exhaustively by
//            tests/test_fastdiv.py), the fp64 sqrt/div of revent.c:69 is replaced by a Newton step on
//            rsqrt whose result is accepted only when it provably rounds to the same float -- otherwise
//            (about 3 positions in 1e5) the exact __dsqrt_rn/__ddiv_rn path runs
//   detect   lane c < k: the state machine over the 32 new t-stat pairs; a peak carries the prefix sum at
//            its position, so events are finished on the spot: raw mean (:160-172), fp64 sum / sum2 in
//            emission order (:163-164).  The last peak never bounds an event (gen_events closes the last
//            event at s_len), hence events trail peaks by one.
//
// Outputs per chunk: raw (un-normalised) events + start samples in a scratch row, event count, fp64 mean
// and standard deviation.  k_scan turns counts into CSR offsets, k_ev_normalise writes the z-normalised
// events (:178-184) to their CSR position.  HBM traffic: the signal once, 8 bytes per event of scratch
// written and read back, 4 (+4 with positions) bytes per event out.
#include "ra_common.cuh"
#include <float.h>

namespace ra {

constexpr int kFusedWarps = 4;
constexpr int kXS = 33;              // row stride of the 32-entry tiles (odd: lane c reading [c][s] is conflict free)
constexpr int kPS = 65;              // row stride of the 64-entry prefix rings
constexpr uint32_t kRingMask = 63;

bool events_fused_supported(uint32_t w1, uint32_t w2) { return w1 >= 1 && w2 >= 1 && w1 <= kFusedMaxWindow && w2 <= kFusedMaxWindow; }
size_t events_fused_smem(int k) { return (size_t)kFusedWarps * k * (2 * kXS + 2 * kPS) * sizeof(float); }

// x / wf, correctly rounded, for wf = a window length in [2, 32] (oracle/ra_oracle.c:ora_check_fastdiv)
__device__ __forceinline__ float fdiv_w(float x, float wf, float rcp)
{
	const float ax = fabsf(x);
	if ((ax >= 1e-30f && ax <= 1e30f) || ax == 0.0f) {
		const float q0 = __fmul_rn(x, rcp);
		const float r = __fmaf_rn(-q0, wf, x);
		return __fmaf_rn(r, rcp, q0);
	}
	return __fdiv_rn(x, wf);
}

// revent.c:50-69 for one window from the prefix values at p-w, p, p+w
template <bool CONTRACT>
__device__ __forceinline__ float tstat_point_fast(float s_m, float s_0, float s_p, float q_m, float q_0, float q_p, float wf, float rcp)
{
	// at p == w the reference skips the subtraction (:53-56); ps[0] = pss[0] = +0.0f so x - 0.0f == x bit for bit
	const float mean1 = fdiv_w(__fsub_rn(s_0, s_m), wf, rcp);
	const float mean2 = fdiv_w(__fsub_rn(s_p, s_0), wf, rcp);
	const float e1 = fdiv_w(__fsub_rn(q_0, q_m), wf, rcp);
	const float e2 = fdiv_w(__fsub_rn(q_p, q_0), wf, rcp);
	float var;
	if (CONTRACT) var = __fmaf_rn(-mean2, mean2, __fadd_rn(__fmaf_rn(-mean1, mean1, e1), e2));            // :61, two vfnmadd
	else var = __fsub_rn(__fadd_rn(__fsub_rn(e1, __fmul_rn(mean1, mean1)), e2), __fmul_rn(mean2, mean2));
	var = fmaxf(var, FLT_MIN);                                                                              // :63
	const float dm = __fsub_rn(mean2, mean1);                                                               // :68
	const float vw = fdiv_w(var, wf, rcp);
	const float x = fabsf(dm);
	// :69  (float)(fabs((double)dm) / sqrt((double)vw)).  y = x * rsqrt(vw) to ~2^-43; the reference's doubly rounded
	// quotient is within 2^-52 of the exact value, so both round to the same float unless y sits within 2^13 double
	// ulps of a float rounding boundary (mantissa bits 28..0 == 0x10000000).
	if (vw >= 1e-30f && vw <= 1e30f && x <= 1e30f) {
		float r0f;
		asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r0f) : "f"(vw));
		const double r0 = (double)r0f, vd = (double)vw;
		const double e = __fma_rn(-__dmul_rn(vd, r0), r0, 1.0);
		const double r1 = __fma_rn(__dmul_rn(r0, 0.5), e, r0);
		const double y = __dmul_rn((double)x, r1);
		const uint32_t lo = (uint32_t)__double2loint(y);
		const bool near_boundary = ((lo - (0x10000000u - 0x2000u)) & 0x1FFFFFFFu) < 0x4000u;
		if (!near_boundary && (y >= 1e-37 || y == 0.0)) return __double2float_rn(y);
	}
	return __double2float_rn(__ddiv_rn((double)x, __dsqrt_rn((double)vw)));
}

struct Det {               // ri_detect_t, revent.c:9-20 (DEF_PEAK_POS = -1, DEF_PEAK_VAL = FLT_MAX)
	uint32_t masked_to;
	int peak_pos;
	float peak_value;
	int valid_peak;
	float peak_ps;         // prefix_sum[peak_pos], captured when peak_pos is assigned (needed by gen_events :162)
};

struct EvAcc {             // gen_events running state (:150-174)
	uint32_t n_peaks;      // peaks emitted so far
	uint32_t start_pos; float start_ps;      // start of the event being built (previous bounded peak, or 0)
	uint32_t pend_pos; float pend_ps;        // last emitted peak: bounds an event only if another peak follows
	double sum, sum2;
	uint32_t n_ev;
};

__device__ __forceinline__ void finish_event(EvAcc &a, uint32_t end_pos, float end_ps, float *raw, uint32_t *raw_pos)
{
	// :162 / :172   (prefix_sum[end] - l_prefixsum) / ((float)end - l_peak)
	const float ev = __fdiv_rn(__fsub_rn(end_ps, a.start_ps), __fsub_rn(__uint2float_rn(end_pos), __uint2float_rn(a.start_pos)));
	a.sum = __dadd_rn(a.sum, (double)ev);                                  // :163
	a.sum2 = __dadd_rn(a.sum2, (double)__fmul_rn(ev, ev));                  // :164 (float product, then widened)
	raw[a.n_ev] = ev;
	if (raw_pos) raw_pos[a.n_ev] = a.start_pos;
	++a.n_ev;
	a.start_pos = end_pos; a.start_ps = end_ps;
}

__device__ __forceinline__ void emit_peak(EvAcc &a, uint32_t pos, float ps_at, float *raw, uint32_t *raw_pos)
{
	if (a.n_peaks > 0) finish_event(a, a.pend_pos, a.pend_ps, raw, raw_pos);
	a.pend_pos = pos; a.pend_ps = ps_at;
	++a.n_peaks;
}

// One detector step, revent.c:89-133.  SHORT selects the masking side effect of :113-120.  ps_i = prefix_sum[i].
template <bool SHORT>
__device__ __forceinline__ void det_step(Det &d, Det &long_d, uint32_t i, float v, float ps_i, float threshold, uint32_t window_length,
                                         float peak_height, EvAcc &acc, float *raw, uint32_t *raw_pos)
{
	if (d.masked_to >= i) return;                                           // :91 (unsigned: i = 0 is always skipped)
	if (d.peak_pos == -1) {                                                 // :94
		if (v < d.peak_value) d.peak_value = v;
		else if (__fsub_rn(v, d.peak_value) > peak_height) { d.peak_value = v; d.peak_pos = (int)i; d.peak_ps = ps_i; }
	} else {
		if (v > d.peak_value) { d.peak_value = v; d.peak_pos = (int)i; d.peak_ps = ps_i; }   // :107-111
		if (SHORT && d.peak_value > threshold) {                            // :113-120
			long_d.masked_to = (uint32_t)d.peak_pos + window_length;
			long_d.peak_pos = -1; long_d.peak_value = FLT_MAX; long_d.valid_peak = 0;
		}
		if (__fsub_rn(d.peak_value, v) > peak_height && d.peak_value > threshold) d.valid_peak = 1;   // :122
		if (d.valid_peak && (i - (uint32_t)d.peak_pos) > window_length / 2) {                            // :126
			emit_peak(acc, (uint32_t)d.peak_pos, d.peak_ps, raw, raw_pos);  // :128
			d.peak_pos = -1; d.peak_value = v; d.valid_peak = 0;
		}
	}
}

struct FusedArgs {
	const float *sig;
	const uint64_t *chunk_off;
	const uint32_t *chunk_len;
	uint32_t n_chunks;
	int k;                       // chunks per warp
	EvParams prm;
	float *raw;                  // [n_chunks][raw_stride] un-normalised events
	uint32_t *raw_pos;           // [n_chunks][raw_stride] first sample of every event (may be null)
	uint32_t raw_stride;
	uint32_t *n_ev;              // events per chunk
	double2 *stats;              // (mean, sd) per chunk
};

constexpr int kFusedMaxK = 16;

template <bool CONTRACT>
__global__ void __launch_bounds__(kFusedWarps * 32)
k_events_fused(FusedArgs A)
{
	extern __shared__ float smem[];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int k = A.k;
	float *xs = smem + (size_t)wid * k * (2 * kXS + 2 * kPS);   // [k][kXS]  samples, then t-stat track 1
	float *t2s = xs + k * kXS;                                    // [k][kXS]  t-stat track 2
	float *psr = t2s + k * kXS;                                   // [k][kPS]  prefix-sum ring
	float *qsr = psr + k * kPS;                                   // [k][kPS]  prefix-sum-of-squares ring
	const uint32_t c0 = (blockIdx.x * kFusedWarps + wid) * (uint32_t)k;
	if (c0 >= A.n_chunks) return;
	const int kk = (int)min((uint32_t)k, A.n_chunks - c0);
	const uint32_t w1 = A.prm.w1, w2 = A.prm.w2;
	const int lag = (int)max(w1, w2);
	const float w1f = (float)w1, w2f = (float)w2, rcp1 = __frcp_rn(w1f), rcp2 = __frcp_rn(w2f);

	// lane c < kk owns chunk c0 + c
	uint32_t my_len = 0; uint64_t my_off = 0;
	if (lane < kk) { my_len = A.chunk_len[c0 + lane]; my_off = A.chunk_off[c0 + lane]; }
	uint32_t max_len = my_len;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(0xffffffffu, max_len, o));
	const int n_tiles = (int)((max_len + (uint32_t)lag + 31u) / 32u);

	float run_s = 0.0f, run_q = 0.0f;                             // ps / pss of the samples consumed so far
	Det sd = { 0u, -1, FLT_MAX, 0, 0.0f }, ld = { 0u, -1, FLT_MAX, 0, 0.0f };   // :198-202
	EvAcc acc = { 0u, 0u, 0.0f, 0u, 0.0f, 0.0, 0.0, 0u };
	float *my_raw = A.raw + (size_t)(c0 + (lane < kk ? lane : 0)) * A.raw_stride;
	uint32_t *my_raw_pos = A.raw_pos ? A.raw_pos + (size_t)(c0 + (lane < kk ? lane : 0)) * A.raw_stride : nullptr;
	if (lane < kk) { psr[lane * kPS] = 0.0f; qsr[lane * kPS] = 0.0f; }   // ps[0] = pss[0] = 0 (:26-27)

	// prefetch tile 0
	float xr[kFusedMaxK];
#pragma unroll
	for (int c = 0; c < kFusedMaxK; ++c) {
		xr[c] = 0.0f;
		if (c < kk) {
			const uint32_t len = __shfl_sync(0xffffffffu, my_len, c);
			const uint64_t off = __shfl_sync(0xffffffffu, my_off, c);
			if ((uint32_t)lane < len) xr[c] = __ldcs(A.sig + off + lane);
		}
	}

	for (int b = 0; b < n_tiles; ++b) {
		// ---- samples of tile b to shared memory; tile b+1 into registers ----
#pragma unroll
		for (int c = 0; c < kFusedMaxK; ++c) {
			if (c < kk) {
				xs[c * kXS + lane] = xr[c];
				const uint32_t len = __shfl_sync(0xffffffffu, my_len, c);
				const uint64_t off = __shfl_sync(0xffffffffu, my_off, c);
				const uint32_t nx = (uint32_t)(b + 1) * 32u + (uint32_t)lane;
				xr[c] = (nx < len) ? __ldcs(A.sig + off + nx) : 0.0f;
			}
		}
		__syncwarp();
		// ---- prefix sums of the 32 new samples (sequential fp32, :28-31) ----
		if (lane < kk) {
			const uint32_t base = (uint32_t)b * 32u;
			const float *xrow = xs + lane * kXS;
			float *prow = psr + lane * kPS, *qrow = qsr + lane * kPS;
#pragma unroll 8
			for (uint32_t s = 0; s < 32u; ++s) {
				if (base + s < my_len) {
					const float x = xrow[s];
					run_s = __fadd_rn(run_s, x);
					run_q = CONTRACT ? __fmaf_rn(x, x, run_q) : __fadd_rn(run_q, __fmul_rn(x, x));
					prow[(base + s + 1u) & kRingMask] = run_s;
					qrow[(base + s + 1u) & kRingMask] = run_q;
				}
			}
		}
		__syncwarp();
		// ---- both t-statistic tracks for positions [32b - lag, 32b - lag + 32) of every chunk ----
		const int p0 = b * 32 - lag;
		for (int c = 0; c < kk; ++c) {
			const int len = (int)__shfl_sync(0xffffffffu, my_len, c);
			const int p = p0 + lane;
			const float *prow = psr + c * kPS, *qrow = qsr + c * kPS;
			float v1 = 0.0f, v2 = 0.0f;
			const bool ok1 = len >= 2 * (int)w1 && w1 >= 2u && p >= (int)w1 && p + (int)w1 <= len;   // :46, :50
			const bool ok2 = len >= 2 * (int)w2 && w2 >= 2u && p >= (int)w2 && p + (int)w2 <= len;
			if (ok1 || ok2) {
				const float s_0 = prow[(uint32_t)p & kRingMask], q_0 = qrow[(uint32_t)p & kRingMask];
				if (ok1) v1 = tstat_point_fast<CONTRACT>(prow[(uint32_t)(p - (int)w1) & kRingMask], s_0, prow[(uint32_t)(p + (int)w1) & kRingMask],
				                                         qrow[(uint32_t)(p - (int)w1) & kRingMask], q_0, qrow[(uint32_t)(p + (int)w1) & kRingMask], w1f, rcp1);
				if (ok2) v2 = tstat_point_fast<CONTRACT>(prow[(uint32_t)(p - (int)w2) & kRingMask], s_0, prow[(uint32_t)(p + (int)w2) & kRingMask],
				                                         qrow[(uint32_t)(p - (int)w2) & kRingMask], q_0, qrow[(uint32_t)(p + (int)w2) & kRingMask], w2f, rcp2);
			}
			xs[c * kXS + lane] = v1; t2s[c * kXS + lane] = v2;
		}
		__syncwarp();
		// ---- the two detectors over the 32 new positions (:87-135), events finished as peaks are emitted ----
		if (lane < kk) {
			const float *t1row = xs + lane * kXS, *t2row = t2s + lane * kXS, *prow = psr + lane * kPS;
			for (int s = 0; s < 32; ++s) {
				const int i = p0 + s;
				if (i >= 0 && i < (int)my_len) {
					const float ps_i = prow[(uint32_t)i & kRingMask];
					det_step<true>(sd, ld, (uint32_t)i, t1row[s], ps_i, A.prm.thr1, w1, A.prm.peak_height, acc, my_raw, my_raw_pos);
					det_step<false>(ld, ld, (uint32_t)i, t2row[s], ps_i, A.prm.thr2, w2, A.prm.peak_height, acc, my_raw, my_raw_pos);
				}
			}
		}
		__syncwarp();
	}
	if (lane < kk) {
		double2 st = make_double2(0.0, 1.0);
		if (acc.n_peaks > 0) {
			finish_event(acc, my_len, run_s, my_raw, my_raw_pos);                  // the last event ends at s_len (:172)
			const double n = (double)acc.n_ev;
			const double mean = __ddiv_rn(acc.sum, n), q = __ddiv_rn(acc.sum2, n); // :179
			const double var = CONTRACT ? __fma_rn(-mean, mean, q) : __dsub_rn(q, __dmul_rn(mean, mean));   // :180
			st = make_double2(mean, __dsqrt_rn(var));
		}
		A.n_ev[c0 + lane] = acc.n_ev;
		A.stats[c0 + lane] = st;
	}
}

// z-normalisation (:182-184) and CSR placement, one warp per chunk.
__global__ void __launch_bounds__(128)
k_ev_normalise(uint32_t n_chunks, const float *__restrict__ raw, const uint32_t *__restrict__ raw_pos, uint32_t raw_stride,
               const uint32_t *__restrict__ n_ev, const double2 *__restrict__ stats, const uint64_t *__restrict__ ev_off,
               float *__restrict__ ev_out, uint64_t ev_cap, uint32_t *__restrict__ ev_pos)
{
	const uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (c >= n_chunks) return;
	const uint32_t n = n_ev[c];
	const double2 st = stats[c];
	const uint64_t off = ev_off[c];
	const float *r = raw + (size_t)c * raw_stride;
	for (uint32_t k = lane; k < n; k += 32)
		if (off + k < ev_cap) {
			ev_out[off + k] = __double2float_rn(__ddiv_rn(__dsub_rn((double)r[k], st.x), st.y));   // :183
			if (ev_pos && raw_pos) ev_pos[off + k] = raw_pos[(size_t)c * raw_stride + k];
		}
}

void events_fused_layout(uint32_t max_len, const EvParams &p, uint32_t *raw_stride, size_t *floats_per_chunk)
{
	// a detector emits at i >= peak_pos + w/2 + 1 and can only start its next peak at a later sample,
	// so peaks of one detector are >= w/2 + 2 apart; events = peaks
	*raw_stride = ((max_len / (p.w1 / 2 + 2) + max_len / (p.w2 / 2 + 2) + 4) + 3u) & ~3u;
	*floats_per_chunk = 2 * (size_t)*raw_stride + 4;             // raw events, their start samples, (mean, sd) as double2
}

cudaError_t launch_events_fused(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                                uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                                uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, int n_sm,
                                cudaStream_t s, cudaEvent_t after_detect)
{
	if (n_chunks == 0) return cudaSuccess;
	uint32_t stride; size_t fpc;
	events_fused_layout(max_len, p, &stride, &fpc);
	FusedArgs A;
	A.sig = sig; A.chunk_off = chunk_off; A.chunk_len = chunk_len; A.n_chunks = n_chunks; A.prm = p;
	A.stats = reinterpret_cast<double2 *>(scratch);               // 16-byte aligned: scratch comes from cudaMalloc
	A.raw = scratch + 4 * (size_t)n_chunks;
	A.raw_pos = reinterpret_cast<uint32_t *>(A.raw + (size_t)n_chunks * stride);
	A.raw_stride = stride; A.n_ev = n_ev;
	// chunks per warp: enough warps to fill the machine (~24 per SM) before lanes are packed
	int k = (int)((n_chunks + (uint32_t)n_sm * 24u - 1u) / ((uint32_t)n_sm * 24u));
	k = k < 1 ? 1 : (k > kFusedMaxK ? kFusedMaxK : k);
	A.k = k;
	const size_t smem = events_fused_smem(k);
	static unsigned long long attr_devices = 0;                       // the opt-in is per function AND per device
	int dev = 0;
	cudaGetDevice(&dev);
	if (!((attr_devices >> (dev & 63)) & 1ull)) {
		cudaError_t ae = cudaFuncSetAttribute(k_events_fused<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)events_fused_smem(kFusedMaxK));
		if (ae == cudaSuccess) ae = cudaFuncSetAttribute(k_events_fused<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)events_fused_smem(kFusedMaxK));
		if (ae != cudaSuccess) return ae;
		__atomic_fetch_or(&attr_devices, 1ull << (dev & 63), __ATOMIC_RELAXED);
	}
	const uint32_t per_cta = (uint32_t)k * kFusedWarps;
	const uint32_t ctas = (n_chunks + per_cta - 1) / per_cta;
	if (contract) k_events_fused<true><<<ctas, kFusedWarps * 32, smem, s>>>(A);
	else k_events_fused<false><<<ctas, kFusedWarps * 32, smem, s>>>(A);
	if (after_detect) cudaEventRecord(after_detect, s);
	cudaError_t e = launch_exclusive_scan_u32_chained(n_ev, ev_off, n_chunks, chained, s);
	if (e != cudaSuccess) return e;
	const unsigned nblocks = (unsigned)(((uint64_t)n_chunks * 32 + 127) / 128);
	k_ev_normalise<<<nblocks, 128, 0, s>>>(n_chunks, A.raw, ev_pos ? A.raw_pos : nullptr, stride, n_ev, A.stats, ev_off, ev_out, ev_cap, ev_pos);
	return cudaGetLastError();
}

} // namespace ra
