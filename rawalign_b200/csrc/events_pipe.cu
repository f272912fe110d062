// events_pipe.cu -- K1 (v4 pipeline, v5 tuning): detect_events() (src/revent.c:190-210) for a batch of chunks as a pipeline of
// kernels that each use exactly the parallelism their stage has.  Bit-identical to the reference build.
//
// What the reference's arithmetic forces to be sequential per chunk: the fp32 prefix sums (revent.c:22-32,
// rounding depends on the order) and the two-detector peak state machine (:77-138).  ncu on the fused v3
// kernel (profiles/r01_ncu_full_v3_summary.md) showed it issue-bound at 7.7 of 32 active lanes because those
// sections ran on k = 3 lanes per warp.  Here every lane of a warp owns one chunk in the sequential stages
// (32 chunks per warp, all lanes busy, nothing but the recurrence on the critical path), HBM latency is
// hidden by a 2-stage cp.async ring of 32-sample x 32-chunk tiles in shared memory, and everything that is
// position-independent runs in fully parallel kernels:
//
//   k_prefix_pipe  lane per chunk   ps/pss recurrences; tiles in by cp.async (16 B, coalesced 128 B rows),
//                                   results out as 128-bit stores.  Row stride 36 floats: 16-byte aligned and
//                                   conflict-free for lane-per-row LDS.128/STS.128.
//   k_tstat_pipe   4 positions/thread both windows (:34-75), 512 positions of a chunk per CTA.  mean2/e2 of position p equal mean1/e1 of position
//                                   p+w (same expression, same operands), so every window mean is divided once
//                                   and shared through shared memory: 3 instead of 5 divisions per window.
//   k_fsm_pipe     lane per chunk   the detector state machine only: peak positions to a per-chunk list.  The
//                                   event arithmetic (:140-188) is hoisted out of the recurrence.
//   k_scan         one CTA          CSR offsets from the per-chunk peak counts (events.cu)
//   k_finish_pipe  warp per chunk   event means in parallel (gather ps[peak]), fp64 sum/sum2 in emission order,
//                                   z-normalisation, CSR write.
//
// HBM traffic per sample: 4 B in + 8 B ps/pss out, 8 B in + 8 B t-stats out, 8 B in = 36 B (scratch is
// L2-absorbed only for small batches); per event 4 B peak + 4 B ps gather + 4-8 B out.
//
// Compiled with -fmad=false: fused multiply-adds exist only where the reference build has them (CONTRACT).
#include "ra_common.cuh"
#include <float.h>
#include <math.h>
#include <stdlib.h>

namespace ra {
namespace evp {

constexpr int kTile = 32;            // samples per tile per chunk
constexpr int kRow = 36;             // shared-memory row stride (floats)
#ifndef RA_PREFIX_STAGES
#define RA_PREFIX_STAGES 2
#endif
#ifndef RA_FSM_STAGES
#define RA_FSM_STAGES 2
#endif
constexpr int kStages = RA_PREFIX_STAGES;      // prefix: input ring depth
constexpr int kFsmStages = RA_FSM_STAGES;     // state machine: ring depth per track
// The lane-per-chunk kernels run kLaneWarps independent warps per CTA, one per SM sub-partition: a CTA's warp i is scheduled
// on sub-partition i % 4, so single-warp CTAs would all pile onto sub-partition 0 (ncu: 2.4 warps on the busiest one with
// two CTAs per SM).  Each warp has a private slice of the dynamic shared memory and synchronises only with itself.
constexpr int kLaneWarps = 4;
constexpr int kTsBlock = 512;
constexpr int kHalo = 16;            // >= kPipeMaxWindow

__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gsrc, int src_bytes)
{
	const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(float *smem_dst, const float *gsrc)
{
	const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Warp-cooperative load of one tile: rows r = 0..31 (chunk c0 + r), columns [base, base + 32) of a per-chunk array,
// into dst[r * kRow + col].  Row r's samples start at rowp[r]; rowl[r] of them exist.  Pieces past the end of a
// row are not loaded (consumers never read past rowl).
template <bool ALIGNED>
__device__ __forceinline__ void load_tile(float *dst, const float *const *rowp, const uint32_t *rowl, uint32_t base, int lane)
{
	const int piece = lane & 7;
	const uint32_t p0 = base + 4u * (uint32_t)piece;
#pragma unroll
	for (int k = 0; k < 8; ++k) {
		const int row = 4 * k + (lane >> 3);
		const uint32_t len = rowl[row];
		if (p0 < len) {
			const float *g = rowp[row] + p0;
			float *d = dst + row * kRow + 4 * piece;
			const uint32_t n = min(4u, len - p0);
			if (ALIGNED || (reinterpret_cast<uintptr_t>(g) & 15u) == 0) cp_async16(d, g, (int)(4u * n));
			else
				for (uint32_t e = 0; e < n; ++e) cp_async4(d + e, g + e);
		}
	}
}

// ---------------------------------------------------------------------------------------------------
// k_prefix_pipe: psx[j] = ps[j+1], pssx[j] = pss[j+1] for j in [0, len)   (ps[0] = pss[0] = 0 is implicit)
// ---------------------------------------------------------------------------------------------------
struct __align__(16) PrefixSmem {
	float in[kStages][32 * kRow];
	float outs[32 * kRow];
	float outq[32 * kRow];
	const float *rowp[32];
	uint32_t rowl[32];
};

template <bool CONTRACT>
__global__ void __launch_bounds__(32 * kLaneWarps)
k_prefix_pipe(const float *__restrict__ sig, const uint64_t *__restrict__ chunk_off, const uint32_t *__restrict__ chunk_len,
              uint32_t n_chunks, float *__restrict__ psx, float *__restrict__ pssx, uint32_t stride)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	PrefixSmem &S = reinterpret_cast<PrefixSmem *>(smem_raw)[threadIdx.x >> 5];
	float (*in)[32 * kRow] = S.in;
	float *outs = S.outs, *outq = S.outq;
	const float **rowp = S.rowp;
	uint32_t *rowl = S.rowl;
	const int lane = threadIdx.x & 31;
	const uint32_t c0 = (blockIdx.x * (uint32_t)kLaneWarps + (threadIdx.x >> 5)) * 32u, c = c0 + (uint32_t)lane;
	if (c0 >= n_chunks) return;                                       // whole warp; warps never synchronise with each other
	const uint32_t my_len = c < n_chunks ? chunk_len[c] : 0u;
	rowp[lane] = c < n_chunks ? sig + chunk_off[c] : sig;
	rowl[lane] = my_len;
	uint32_t max_len = my_len;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(0xffffffffu, max_len, o));
	__syncwarp();
	const int n_tiles = (int)((max_len + kTile - 1) / kTile);
#pragma unroll
	for (int t = 0; t < kStages - 1; ++t) {
		if (t < n_tiles) load_tile<false>(in[t], rowp, rowl, (uint32_t)t * kTile, lane);
		cp_async_commit();
	}
	float s = 0.0f, q = 0.0f;                                         // revent.c:26-27
	const int piece = lane & 7;
	for (int t = 0; t < n_tiles; ++t) {
		if (t + kStages - 1 < n_tiles) load_tile<false>(in[(t + kStages - 1) % kStages], rowp, rowl, (uint32_t)(t + kStages - 1) * kTile, lane);
		cp_async_commit();
		cp_async_wait<kStages - 1>();
		__syncwarp();
		const uint32_t base = (uint32_t)t * kTile;
		if (base < my_len) {
			const float *row = in[t % kStages] + lane * kRow;
			const uint32_t cnt = min((uint32_t)kTile, my_len - base);
#pragma unroll
			for (int k = 0; k < 8; ++k) {
				const float4 v = *reinterpret_cast<const float4 *>(row + 4 * k);
				const float x[4] = { v.x, v.y, v.z, v.w };
				float a[4], b[4];
#pragma unroll
				for (int e = 0; e < 4; ++e) {
					if ((uint32_t)(4 * k + e) < cnt) {
						s = __fadd_rn(s, x[e]);                                                                  // :29
						q = CONTRACT ? __fmaf_rn(x[e], x[e], q) : __fadd_rn(q, __fmul_rn(x[e], x[e]));             // :30
					}
					a[e] = s; b[e] = q;
				}
				*reinterpret_cast<float4 *>(outs + lane * kRow + 4 * k) = make_float4(a[0], a[1], a[2], a[3]);
				*reinterpret_cast<float4 *>(outq + lane * kRow + 4 * k) = make_float4(b[0], b[1], b[2], b[3]);
			}
		}
		__syncwarp();
		const uint32_t p0 = base + 4u * (uint32_t)piece;
#pragma unroll
		for (int k = 0; k < 8; ++k) {
			const int row = 4 * k + (lane >> 3);
			if (p0 < rowl[row]) {                                      // whole 16-byte piece: rows are padded to the stride
				const size_t o = (size_t)(c0 + (uint32_t)row) * stride + p0;
				*reinterpret_cast<float4 *>(psx + o) = *reinterpret_cast<const float4 *>(outs + row * kRow + 4 * piece);
				*reinterpret_cast<float4 *>(pssx + o) = *reinterpret_cast<const float4 *>(outq + row * kRow + 4 * piece);
			}
		}
		__syncwarp();
	}
}

// ---------------------------------------------------------------------------------------------------
// t-statistics
// ---------------------------------------------------------------------------------------------------
// x / wf, correctly rounded, for wf = a window length in [2, 32] and 1e-30 <= |x| <= 1e30 (exhaustive check of the
// three-instruction sequence: tests/test_fastdiv.py); anything else (zeros, subnormals, huge values) takes the IEEE division.
__device__ __forceinline__ float fdiv_w(float x, float wf, float rcp)
{
	const float ax = fabsf(x);
	if (ax >= 1e-30f && ax <= 1e30f) {
		const float q0 = __fmul_rn(x, rcp);
		const float r = __fmaf_rn(-q0, wf, x);
		return __fmaf_rn(r, rcp, q0);
	}
	return __fdiv_rn(x, wf);
}

// Per-window constants of the clamped-variance case: with fp32 prefix sums of ~1e7 the window variance of revent.c:61 is
// rounding noise late in a chunk and is clamped to FLT_MIN at :63 for a sizeable share of positions (ncu: 19 % of warps had
// such a lane), so that case gets a straight-line path of its own instead of the IEEE fallbacks.
struct TsWin {
	float wf, rcp;
	float vw_min;        // FLT_MIN / wf            (fp32 division, :69's var/w)
	double rs_min;       // 1 / sqrt((double)vw_min)
};

// revent.c:61-69 from the window means (mean1, mean2) and mean squares (e1, e2)
template <bool CONTRACT>
__device__ __forceinline__ float tstat_from_means(float mean1, float mean2, float e1, float e2, const TsWin W)
{
	float var;
	if (CONTRACT) var = __fmaf_rn(-mean2, mean2, __fadd_rn(__fmaf_rn(-mean1, mean1, e1), e2));            // :61, two vfnmadd
	else var = __fsub_rn(__fadd_rn(__fsub_rn(e1, __fmul_rn(mean1, mean1)), e2), __fmul_rn(mean2, mean2));
	const bool clamped = !(var > FLT_MIN);                                                                  // :63 fmaxf(var, FLT_MIN); NaN -> FLT_MIN too
	const float dm = __fsub_rn(mean2, mean1);                                                               // :68
	const float x = fabsf(dm);
	const float vw = clamped ? W.vw_min : fdiv_w(var, W.wf, W.rcp);
	// :69  (float)(fabs((double)dm) / sqrt((double)vw)).  y = x * r with r = 1/sqrt(vw) to ~2^-50 (one Newton step in fp64
	// from the fp32 rsqrt; the host-computed constant when clamped); the reference's doubly rounded quotient is within 2^-52
	// of the exact value, so both round to the same float unless y sits within 2^13 double ulps of a float rounding
	// boundary -- then, and for operands outside the comfortable range, the exact path runs.
	const bool ok = clamped || (vw >= 1e-30f && vw <= 1e30f);
	if (ok && x <= 1e30f) {
		const float vs = clamped ? 1.0f : vw;
		float r0f;
		asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r0f) : "f"(vs));
		const double r0 = (double)r0f, vd = (double)vs;
		const double e = __fma_rn(-__dmul_rn(vd, r0), r0, 1.0);
		const double rn = __fma_rn(__dmul_rn(r0, 0.5), e, r0);
		const double r1 = clamped ? W.rs_min : rn;
		const double y = __dmul_rn((double)x, r1);
		const uint32_t lo = (uint32_t)__double2loint(y);
		const bool near_boundary = ((lo - (0x10000000u - 0x2000u)) & 0x1FFFFFFFu) < 0x4000u;
		if (!near_boundary && (y >= 1e-37 || y == 0.0)) return __double2float_rn(y);
	}
	return __double2float_rn(__ddiv_rn((double)x, __dsqrt_rn((double)vw)));
}

// out-of-line copies for the rare cases of k_tstat_pipe (keeps its hot loop compact)
template <bool CONTRACT>
__device__ __noinline__ float tstat_exact(float mean1, float mean2, float e1, float e2, float wf, float rcp, float vw_min, double rs_min)
{
	const TsWin W = { wf, rcp, vw_min, rs_min };
	return tstat_from_means<CONTRACT>(mean1, mean2, e1, e2, W);
}
__device__ __noinline__ float fdiv_w_exact(float x, float wf, float rcp) { return fdiv_w(x, wf, rcp); }

// The common case of tstat_from_means as straight-line code: the same operations in the same order, with every range
// condition folded into one flag instead of a branch per operation.  ok == false means "take tstat_from_means".
template <bool CONTRACT>
__device__ __forceinline__ float tstat_fast(float mean1, float mean2, float e1, float e2, const TsWin W, bool &ok)
{
	float var;
	if (CONTRACT) var = __fmaf_rn(-mean2, mean2, __fadd_rn(__fmaf_rn(-mean1, mean1, e1), e2));            // :61
	else var = __fsub_rn(__fadd_rn(__fsub_rn(e1, __fmul_rn(mean1, mean1)), e2), __fmul_rn(mean2, mean2));
	const bool clamped = !(var > FLT_MIN);                                                                  // :63
	const float x = fabsf(__fsub_rn(mean2, mean1));                                                         // :68
	const float q0 = __fmul_rn(var, W.rcp);                                                                 // fdiv_w(var)
	const float vwf = __fmaf_rn(__fmaf_rn(-q0, W.wf, var), W.rcp, q0);
	const float vs = clamped ? 1.0f : vwf;
	float r0f;
	asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r0f) : "f"(vs));
	const double r0 = (double)r0f, vd = (double)vs;
	const double e = __fma_rn(-__dmul_rn(vd, r0), r0, 1.0);
	const double rn = __fma_rn(__dmul_rn(r0, 0.5), e, r0);
	const double r1 = clamped ? W.rs_min : rn;
	const double y = __dmul_rn((double)x, r1);
	const uint32_t lo = (uint32_t)__double2loint(y);
	const bool near_boundary = ((lo - (0x10000000u - 0x2000u)) & 0x1FFFFFFFu) < 0x4000u;
	const bool in_range = clamped | ((var <= 1e30f) & (var >= 1e-30f) & (vwf >= 1e-30f) & (vwf <= 1e30f));
	ok = in_range & (x <= 1e30f) & !near_boundary & ((y >= 1e-37) | (y == 0.0));
	return __double2float_rn(y);
}

// grid: x = tiles of kTsBlock positions, y = chunk.  Thread u owns positions base + u + kTsThreads * e, e < kTsPer
// (conflict-free shared memory, coalesced stores); the window means of a position are computed once and shared.
constexpr int kTsThreads = 128;
constexpr int kTsPer = kTsBlock / kTsThreads;
constexpr int kTsLead = 20;          // floats staged before psx[base]: >= kPipeMaxWindow + 1, multiple of 4
constexpr int kTsSp = kTsLead + kTsBlock + kHalo;           // ps[j] sits at index j - base + kTsLead - 1
constexpr int kTsMeans = kTsBlock + kHalo;
static_assert(kTsLead % 4 == 0 && kTsLead >= (int)kPipeMaxWindow + 1 && kHalo >= (int)kPipeMaxWindow && kHalo % 4 == 0 && kHalo <= kTsThreads, "t-stat tile");
template <bool CONTRACT>
__global__ void __launch_bounds__(kTsThreads)
k_tstat_pipe(const uint32_t *__restrict__ chunk_len, const float *__restrict__ psx, const float *__restrict__ pssx, uint32_t stride,
             float *__restrict__ t1, float *__restrict__ t2, uint32_t w1, uint32_t w2, TsWin W1, TsWin W2)
{
	__shared__ __align__(16) float sp[kTsSp], sq[kTsSp];
	__shared__ float M1[kTsMeans], E1[kTsMeans], M2[kTsMeans], E2[kTsMeans];
	const uint32_t c = blockIdx.y;
	const uint32_t len = chunk_len[c];
	const uint32_t base = blockIdx.x * kTsBlock;
	if (base >= len) return;
	const uint32_t u = threadIdx.x;
	{
		// 16-byte pieces f of psx[base - kTsLead + 4f ...]; psx[-1] stands for ps[0] = +0.0f (:26-27), nothing before it is used.
		// Pieces past the chunk's samples hold stale scratch (never consumed: every use below is guarded by len).
		const float4 *__restrict__ prow = reinterpret_cast<const float4 *>(psx + (size_t)c * stride + base) - kTsLead / 4;
		const float4 *__restrict__ qrow = reinterpret_cast<const float4 *>(pssx + (size_t)c * stride + base) - kTsLead / 4;
		const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
		for (uint32_t f = u; f < (uint32_t)(kTsSp / 4); f += kTsThreads) {
			const bool in = base != 0u || f >= (uint32_t)(kTsLead / 4);
			reinterpret_cast<float4 *>(sp)[f] = in ? __ldg(prow + f) : z;
			reinterpret_cast<float4 *>(sq)[f] = in ? __ldg(qrow + f) : z;
		}
	}
	__syncthreads();
	// window mean ending at j = base + m:  M(j) = (ps[j] - ps[j-w]) / w, meaningful for j in [w, len].  At j == w the
	// reference skips the subtraction (:53-56); ps[0] = +0.0f and x - 0.0f == x bit for bit.  Computed for every m
	// (garbage in, garbage out where j is outside that range -- nothing reads those).
	float m1o[kTsPer], e1o[kTsPer], m2o[kTsPer], e2o[kTsPer];
	auto means = [&](uint32_t m, float &a1, float &b1, float &a2, float &b2) {
		const uint32_t k = m + (uint32_t)(kTsLead - 1);
		const float s0 = sp[k], q0 = sq[k];
		const float x1 = __fsub_rn(s0, sp[k - w1]), y1 = __fsub_rn(q0, sq[k - w1]);
		const float x2 = __fsub_rn(s0, sp[k - w2]), y2 = __fsub_rn(q0, sq[k - w2]);
		const float f1 = __fmul_rn(x1, W1.rcp), g1 = __fmul_rn(y1, W1.rcp), f2 = __fmul_rn(x2, W2.rcp), g2 = __fmul_rn(y2, W2.rcp);
		a1 = __fmaf_rn(__fmaf_rn(-f1, W1.wf, x1), W1.rcp, f1);
		b1 = __fmaf_rn(__fmaf_rn(-g1, W1.wf, y1), W1.rcp, g1);
		a2 = __fmaf_rn(__fmaf_rn(-f2, W2.wf, x2), W2.rcp, f2);
		b2 = __fmaf_rn(__fmaf_rn(-g2, W2.wf, y2), W2.rcp, g2);
		const float lo = fminf(fminf(fabsf(x1), fabsf(y1)), fminf(fabsf(x2), fabsf(y2)));
		const float hi = fmaxf(fmaxf(fabsf(x1), fabsf(y1)), fmaxf(fabsf(x2), fabsf(y2)));
		const uint32_t j = base + m;
		if (!(lo >= 1e-30f && hi <= 1e30f) && j <= len) {             // rare: zeros, subnormals, huge or NaN sums -> IEEE division
			if (j >= w1) { a1 = fdiv_w_exact(x1, W1.wf, W1.rcp); b1 = fdiv_w_exact(y1, W1.wf, W1.rcp); }
			if (j >= w2) { a2 = fdiv_w_exact(x2, W2.wf, W2.rcp); b2 = fdiv_w_exact(y2, W2.wf, W2.rcp); }
		}
		M1[m] = a1; E1[m] = b1; M2[m] = a2; E2[m] = b2;
	};
#pragma unroll
	for (int e = 0; e < kTsPer; ++e) means(u + (uint32_t)(kTsThreads * e), m1o[e], e1o[e], m2o[e], e2o[e]);
	if (u < (uint32_t)kHalo) { float a, b, cc, d; means((uint32_t)kTsBlock + u, a, b, cc, d); }
	__syncthreads();
	// :46, :50  valid iff w >= 2, len >= 2w, w <= p <= len - w   (one unsigned compare once len >= 2w)
	const bool any1 = w1 >= 2u && len >= 2u * w1, any2 = w2 >= 2u && len >= 2u * w2;
	const uint32_t span1 = len - 2u * w1, span2 = len - 2u * w2;
	float *__restrict__ o1 = t1 + (size_t)c * stride + base;
	float *__restrict__ o2 = t2 + (size_t)c * stride + base;
#pragma unroll
	for (int e = 0; e < kTsPer; ++e) {
		const uint32_t m = u + (uint32_t)(kTsThreads * e), p = base + m;
		const bool valid1 = any1 && p - w1 <= span1, valid2 = any2 && p - w2 <= span2;
		const float m1n = M1[m + w1], e1n = E1[m + w1], m2n = M2[m + w2], e2n = E2[m + w2];
		bool ok1, ok2;
		float v1 = tstat_fast<CONTRACT>(m1o[e], m1n, e1o[e], e1n, W1, ok1);
		float v2 = tstat_fast<CONTRACT>(m2o[e], m2n, e2o[e], e2n, W2, ok2);
		if (valid1 && !ok1) v1 = tstat_exact<CONTRACT>(m1o[e], m1n, e1o[e], e1n, W1.wf, W1.rcp, W1.vw_min, W1.rs_min);
		if (valid2 && !ok2) v2 = tstat_exact<CONTRACT>(m2o[e], m2n, e2o[e], e2n, W2.wf, W2.rcp, W2.vw_min, W2.rs_min);
		if (p < len) {
			o1[m] = valid1 ? v1 : 0.0f;
			o2[m] = valid2 ? v2 : 0.0f;
		}
	}
}

// ---------------------------------------------------------------------------------------------------
// k_fsm_pipe: gen_peaks, revent.c:77-138
// ---------------------------------------------------------------------------------------------------
// One step of both detectors, revent.c:89-133, as straight-line predicated code (the lanes of a warp are in different
// states at every step, so branches would serialise every path anyway; ncu on the branchy form: 116 warp
// instructions per step at 14.5 active lanes).  Semantics, per detector, in the reference's order:
//   masked_to >= i -> skip;  searching (pos == -1): v < pv -> pv = v, else v - pv > h -> pv = v, pos = i;
//   tracking: v > pv -> pv = v, pos = i;  [short only] pv > thr -> mask + reset the long detector (:113-120);
//   pv - v > h && pv > thr -> valid;  valid && i - pos > w/2 -> emit pos, pos = -1, pv = v, valid = 0.
// The short detector is never masked (only the long one is, :116), so its masked_to stays 0: active iff i > 0.
struct DetS { int pos; float pv; unsigned valid; };
struct DetL { uint32_t masked_to; int pos; float pv; unsigned valid; };

__device__ __forceinline__ void fsm_step(DetS &S, DetL &L, uint32_t i, float v1, float v2, float thr1, float thr2, float h,
                                         uint32_t w1, uint32_t half1, uint32_t half2, uint32_t *__restrict__ peaks, uint32_t &n,
                                         unsigned out_on = 1u)
{
	// bitwise (&, |) on 0/1 flags instead of && / ||: no short-circuit evaluation, so nothing here needs a branch
	// ---- short detector ----
	{
		const unsigned act = i > 0u;
		const unsigned srch = S.pos < 0;
		const unsigned lt = v1 < S.pv;
		const unsigned rise = __fsub_rn(v1, S.pv) > h;
		const unsigned gt = v1 > S.pv;
		const unsigned setpos = act & ((srch & (lt ^ 1u) & rise) | ((srch ^ 1u) & gt));
		const unsigned upd = act & ((srch & (lt | rise)) | ((srch ^ 1u) & gt));
		const float pv = upd ? v1 : S.pv;
		const int pos = setpos ? (int)i : S.pos;
		const unsigned trk = act & (srch ^ 1u);
		const unsigned over = pv > thr1;
		const unsigned dom = trk & over;                                // :113-120
		L.masked_to = dom ? (uint32_t)pos + w1 : L.masked_to;
		L.pos = dom ? -1 : L.pos;
		L.pv = dom ? FLT_MAX : L.pv;
		L.valid = dom ? 0u : L.valid;
		const unsigned valid = S.valid | (dom & (unsigned)(__fsub_rn(pv, v1) > h));     // :122
		const unsigned emit = trk & valid & (unsigned)((i - (uint32_t)pos) > half1);    // :126
		if (emit & out_on) peaks[n] = (uint32_t)pos;                 // :128
		n += emit & out_on;
		S.pos = emit ? -1 : pos; S.pv = emit ? v1 : pv; S.valid = emit ? 0u : valid;
	}
	// ---- long detector ----
	{
		const unsigned act = L.masked_to < i;
		const unsigned srch = L.pos < 0;
		const unsigned lt = v2 < L.pv;
		const unsigned rise = __fsub_rn(v2, L.pv) > h;
		const unsigned gt = v2 > L.pv;
		const unsigned setpos = act & ((srch & (lt ^ 1u) & rise) | ((srch ^ 1u) & gt));
		const unsigned upd = act & ((srch & (lt | rise)) | ((srch ^ 1u) & gt));
		const float pv = upd ? v2 : L.pv;
		const int pos = setpos ? (int)i : L.pos;
		const unsigned trk = act & (srch ^ 1u);
		const unsigned valid = L.valid | (trk & (unsigned)(pv > thr2) & (unsigned)(__fsub_rn(pv, v2) > h));
		const unsigned emit = trk & valid & (unsigned)((i - (uint32_t)pos) > half2);
		if (emit & out_on) peaks[n] = (uint32_t)pos;
		n += emit & out_on;
		L.pos = emit ? -1 : pos; L.pv = emit ? v2 : pv; L.valid = emit ? 0u : valid;
	}
}

struct __align__(16) FsmSmem {
	float s1[kFsmStages][32 * kRow];
	float s2[kFsmStages][32 * kRow];
	const float *rowp1[32];
	const float *rowp2[32];
	uint32_t rowl[32];
};

// SEG > 1: every chunk is cut into SEG segments at tile boundaries and each segment gets its own lane (32 / SEG chunks per
// warp), which divides the length of the recurrence -- the only thing this kernel is bound by -- by SEG.  A segment after
// the first cannot know the detectors' state at its first sample, so it starts kFsmWarm samples early from the reset
// state with emission switched off; the detectors re-synchronise at the next emitted peak (every ~9 samples), so by the
// segment's first sample its state is, in practice always, the true one.  That is VERIFIED, not assumed: the state a
// segment starts its own samples with must equal, field by field and bit for bit, the state the previous segment ends
// with (equal state => identical continuation, so the concatenated peak lists are exactly the sequential ones).  A
// chunk with any mismatch is replayed sequentially by its first lane (test hook: force_replay).  Peaks are partitioned
// by EMISSION step, so a peak found late in one segment and emitted early in the next belongs to the next one.
// A masked_to below the junction is equivalent to 0 (act = masked_to < i holds either way from there on).
constexpr int kFsmMaxSeg = 4;
constexpr uint32_t kFsmWarm = 64;        // warm-up samples (2 tiles); must be a multiple of kTile
constexpr uint32_t kFsmMinSplit = 512;   // shorter chunks run unsplit on their first lane

template <int SEG>
__global__ void __launch_bounds__(32 * kLaneWarps)
k_fsm_pipe(const uint32_t *__restrict__ chunk_len, uint32_t n_chunks, const float *__restrict__ t1, const float *__restrict__ t2,
           uint32_t stride, EvParams prm, uint32_t *__restrict__ peaks_base, uint32_t peak_stride, uint32_t *__restrict__ n_peaks_out,
           uint32_t *__restrict__ seg_counts, int force_replay)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	FsmSmem &S = reinterpret_cast<FsmSmem *>(smem_raw)[threadIdx.x >> 5];
	float (*s1)[32 * kRow] = S.s1;
	float (*s2)[32 * kRow] = S.s2;
	const float **rowp1 = S.rowp1, **rowp2 = S.rowp2;
	uint32_t *rowl = S.rowl;
	constexpr int kStages = kFsmStages;
	constexpr uint32_t kFull = 0xffffffffu;
	constexpr int CPW = 32 / SEG;                                      // chunks per warp
	const int lane = threadIdx.x & 31;
	const int seg = lane % SEG, ci = lane / SEG;                       // the SEG lanes of a chunk are adjacent
	const uint32_t c0 = (blockIdx.x * (uint32_t)kLaneWarps + (threadIdx.x >> 5)) * (uint32_t)CPW, c = c0 + (uint32_t)ci;
	if (c0 >= n_chunks) return;                                       // whole warp; warps never synchronise with each other
	const uint32_t cc = min(c, n_chunks - 1);
	const uint32_t len = c < n_chunks ? chunk_len[c] : 0u;
	const uint32_t segcap = peak_stride / (uint32_t)SEG;
	const float thr1 = prm.thr1, thr2 = prm.thr2, ph = prm.peak_height;
	const uint32_t w1 = prm.w1, half1 = prm.w1 / 2, half2 = prm.w2 / 2;
	// segment [b0, b1) of the chunk; the lane starts at ws <= b0
	const bool split = SEG > 1 && len >= kFsmMinSplit;
	uint32_t b0 = seg == 0 ? 0u : len, b1 = len, ws = b0;
	if (split) {
		const uint32_t q = (((len + SEG - 1) / SEG) + kTile - 1) / kTile * kTile;   // >= 128 > kFsmWarm
		b0 = min(len, (uint32_t)seg * q); b1 = min(len, (uint32_t)(seg + 1) * q);
		ws = (seg > 0 && b0 < b1) ? b0 - kFsmWarm : b0;
	}
	const bool warm = ws != b0;

	DetS sd; DetL ld; uint32_t n_peaks;
	DetS snap_s = { -1, FLT_MAX, 0u }; DetL snap_l = { 0u, -1, FLT_MAX, 0u };
	// runs the lane over samples [from, to) of its chunk; emission on from sample `emit_from`; lanes with from >= to idle
	auto run = [&](uint32_t from, uint32_t to, uint32_t emit_from, uint32_t *__restrict__ peaks, bool take_snapshot) {
		const uint32_t rl = to > from ? to - from : 0u;
		__syncwarp();
		rowp1[lane] = t1 + (size_t)cc * stride + from;
		rowp2[lane] = t2 + (size_t)cc * stride + from;
		rowl[lane] = rl;
		uint32_t max_len = rl;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(kFull, max_len, o));
		__syncwarp();
		const int n_tiles = (int)((max_len + kTile - 1) / kTile);
#pragma unroll
		for (int t = 0; t < kStages - 1; ++t) {
			if (t < n_tiles) { load_tile<true>(s1[t], rowp1, rowl, (uint32_t)t * kTile, lane); load_tile<true>(s2[t], rowp2, rowl, (uint32_t)t * kTile, lane); }
			cp_async_commit();
		}
		for (int t = 0; t < n_tiles; ++t) {
			if (t + kStages - 1 < n_tiles) {
				const int st = (t + kStages - 1) % kStages;
				load_tile<true>(s1[st], rowp1, rowl, (uint32_t)(t + kStages - 1) * kTile, lane);
				load_tile<true>(s2[st], rowp2, rowl, (uint32_t)(t + kStages - 1) * kTile, lane);
			}
			cp_async_commit();
			cp_async_wait<kStages - 1>();
			__syncwarp();
			const uint32_t base = (uint32_t)t * kTile;
			if (base < rl) {
				const float *r1 = s1[t % kStages] + lane * kRow, *r2 = s2[t % kStages] + lane * kRow;
				const uint32_t cnt = min((uint32_t)kTile, rl - base);
				const uint32_t i0 = from + base;
				const unsigned on = SEG == 1 ? 1u : (unsigned)(i0 >= emit_from);   // emit_from is a tile boundary of this lane
				if (cnt == (uint32_t)kTile) {                           // full tile: no per-step guard
#pragma unroll 2
					for (int k = 0; k < 8; ++k) {
						const float4 a = *reinterpret_cast<const float4 *>(r1 + 4 * k);
						const float4 b = *reinterpret_cast<const float4 *>(r2 + 4 * k);
						const uint32_t i = i0 + (uint32_t)(4 * k);
						fsm_step(sd, ld, i, a.x, b.x, thr1, thr2, ph, w1, half1, half2, peaks, n_peaks, on);
						fsm_step(sd, ld, i + 1u, a.y, b.y, thr1, thr2, ph, w1, half1, half2, peaks, n_peaks, on);
						fsm_step(sd, ld, i + 2u, a.z, b.z, thr1, thr2, ph, w1, half1, half2, peaks, n_peaks, on);
						fsm_step(sd, ld, i + 3u, a.w, b.w, thr1, thr2, ph, w1, half1, half2, peaks, n_peaks, on);
					}
				} else {
#pragma unroll 1
					for (uint32_t e = 0; e < cnt; ++e) fsm_step(sd, ld, i0 + e, r1[e], r2[e], thr1, thr2, ph, w1, half1, half2, peaks, n_peaks, on);
				}
				if (SEG > 1 && take_snapshot && i0 + (uint32_t)kTile == emit_from) { snap_s = sd; snap_l = ld; }   // state before the segment's own first sample
			}
			__syncwarp();
		}
	};

	uint32_t *__restrict__ peaks_c = peaks_base + (size_t)cc * peak_stride;
	sd = { -1, FLT_MAX, 0u }; ld = { 0u, -1, FLT_MAX, 0u }; n_peaks = 0;   // :198-202
	run(ws, b1, b0, peaks_c + (size_t)seg * segcap, warm);

	bool replay = false;
	if (SEG > 1) {
		// the state this lane ends with against the snapshot of the next segment's lane
		const int nxt = min(lane + 1, 31);
		const unsigned n_warm = __shfl_sync(kFull, (unsigned)warm, nxt);
		const int ns_pos = __shfl_sync(kFull, snap_s.pos, nxt); const float ns_pv = __shfl_sync(kFull, snap_s.pv, nxt);
		const unsigned ns_valid = __shfl_sync(kFull, snap_s.valid, nxt);
		const uint32_t nl_mask = __shfl_sync(kFull, snap_l.masked_to, nxt); const int nl_pos = __shfl_sync(kFull, snap_l.pos, nxt);
		const float nl_pv = __shfl_sync(kFull, snap_l.pv, nxt); const unsigned nl_valid = __shfl_sync(kFull, snap_l.valid, nxt);
		bool ok = true;
		if (seg + 1 < SEG && n_warm) {                                  // the next lane exists, belongs to this chunk and speculated
			const uint32_t ma = ld.masked_to < b1 ? 0u : ld.masked_to, mb = nl_mask < b1 ? 0u : nl_mask;
			ok = sd.pos == ns_pos && __float_as_uint(sd.pv) == __float_as_uint(ns_pv) && sd.valid == ns_valid && ma == mb &&
			     ld.pos == nl_pos && __float_as_uint(ld.pv) == __float_as_uint(nl_pv) && ld.valid == nl_valid;
		}
		if (force_replay && split) ok = false;
		const unsigned bad = __ballot_sync(kFull, !ok);
		replay = ((bad >> (ci * SEG)) & ((1u << SEG) - 1u)) != 0u;
		if (bad) {                                                     // some chunk of the warp is replayed by its first lane
			const bool mine = replay && seg == 0;
			if (mine) { sd = { -1, FLT_MAX, 0u }; ld = { 0u, -1, FLT_MAX, 0u }; n_peaks = 0; }
			run(0u, mine ? len : 0u, 0u, peaks_c, false);
			if (replay && seg != 0) n_peaks = 0;
		}
	}
	uint32_t tot = n_peaks;
#pragma unroll
	for (int o = 1; o < SEG; o <<= 1) tot += __shfl_xor_sync(kFull, tot, o);   // the SEG lanes of a chunk are an aligned group
	if (c < n_chunks) {
		if (SEG > 1) seg_counts[(size_t)c * SEG + seg] = n_peaks;
		if (seg == 0) n_peaks_out[c] = tot;                            // == number of events (see k_finish_pipe)
	}
}

// ---------------------------------------------------------------------------------------------------
// k_finish_pipe: gen_events, revent.c:140-188.  One warp per chunk.
//
// gen_events consumes peaks[0 .. n_ev-2] in emission order and closes the last event at s_len (:160-172): the final
// peak of a chunk never bounds an event.  n_ev == n_peaks because the filter of :145-147 (0 < peaks[i] < s_len)
// always holds: position 0 is never examined (masked_to(0) >= 0, :91) and positions are < s_len by construction.
// ---------------------------------------------------------------------------------------------------
template <bool CONTRACT, bool SEGS>
__global__ void __launch_bounds__(128)
k_finish_pipe(const uint32_t *__restrict__ chunk_len, uint32_t n_chunks, const float *__restrict__ psx, uint32_t stride,
              const uint32_t *__restrict__ peaks_base, uint32_t peak_stride, const uint32_t *__restrict__ n_peaks_in,
              const uint64_t *__restrict__ ev_off, float *__restrict__ ev_out, uint64_t ev_cap, float *__restrict__ raw_base,
              uint32_t *__restrict__ ev_pos, const uint32_t *__restrict__ seg_counts, int n_seg)
{
	const uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (c >= n_chunks) return;
	const uint32_t n_ev = n_peaks_in[c];
	if (n_ev == 0) return;                                            // detect_events returns NULL, *n untouched (:205-206)
	const uint32_t len = chunk_len[c];
	const float *__restrict__ ps1 = psx + (size_t)c * stride;         // ps[j] = ps1[j-1], j >= 1
	const uint32_t *__restrict__ peaks_c = peaks_base + (size_t)c * peak_stride;
	float *__restrict__ raw = raw_base + (size_t)c * peak_stride;
	// the state machine leaves the chunk's peaks as up to kFsmMaxSeg lists (one per segment, in order): k-th peak overall
	uint32_t cum[kFsmMaxSeg];
	{
		uint32_t acc = 0;
#pragma unroll
		for (int q = 0; q < kFsmMaxSeg; ++q) { acc += (SEGS && q < n_seg) ? seg_counts[(size_t)c * n_seg + q] : 0u; cum[q] = acc; }
	}
	const uint32_t segcap = peak_stride / (uint32_t)max(n_seg, 1);
	auto peak_at = [&](uint32_t k) -> uint32_t {
		if (!SEGS) return peaks_c[k];
		uint32_t sgm = 0, base = 0;
#pragma unroll
		for (int q = 0; q + 1 < kFsmMaxSeg; ++q) { const bool past = k >= cum[q] && q + 1 < n_seg; sgm += past ? 1u : 0u; base = past ? cum[q] : base; }
		return peaks_c[sgm * segcap + (k - base)];
	};
	// event k: (ps[end_k] - ps[start_k]) / ((float)end_k - (float)start_k), end_k = peaks[k] (k < n_ev-1) or len,
	// start_k = peaks[k-1] or 0                                      :160-172
	double sum = 0.0, sum2 = 0.0;
	for (uint32_t kb = 0; kb < n_ev; kb += 32) {
		const uint32_t k = kb + (uint32_t)lane;
		float ev = 0.0f;
		if (k < n_ev) {
			const uint32_t e_pos = (k + 1 < n_ev) ? peak_at(k) : len;
			const uint32_t s_pos = (k > 0) ? peak_at(k - 1) : 0u;
			const float l_ps = (s_pos > 0) ? __ldg(ps1 + s_pos - 1) : 0.0f;
			const float e_ps = (e_pos > 0) ? __ldg(ps1 + e_pos - 1) : 0.0f;
			ev = __fdiv_rn(__fsub_rn(e_ps, l_ps), __fsub_rn(__uint2float_rn(e_pos), __uint2float_rn(s_pos)));
			raw[k] = ev;
		}
		// sum / sum2 in array order, fp64 (:163-164, :173-174): every lane carries the same accumulators
		const uint32_t cnt = min(32u, n_ev - kb);
		if (cnt == 32u) {
#pragma unroll
			for (int t = 0; t < 32; ++t) {
				const float e = __shfl_sync(0xffffffffu, ev, t);
				sum = __dadd_rn(sum, (double)e);
				sum2 = __dadd_rn(sum2, (double)__fmul_rn(e, e));       // float product, then widened (:164)
			}
		} else {
			for (uint32_t t = 0; t < cnt; ++t) {
				const float e = __shfl_sync(0xffffffffu, ev, (int)t);
				sum = __dadd_rn(sum, (double)e);
				sum2 = __dadd_rn(sum2, (double)__fmul_rn(e, e));
			}
		}
	}
	const double mean = __ddiv_rn(sum, (double)n_ev);                  // :179
	const double q = __ddiv_rn(sum2, (double)n_ev);
	const double var = CONTRACT ? __fma_rn(-mean, mean, q) : __dsub_rn(q, __dmul_rn(mean, mean));   // :180
	const double sd = __dsqrt_rn(var);
	const uint64_t off = ev_off[c];
	__syncwarp();
	for (uint32_t k = lane; k < n_ev; k += 32)
		if (off + k < ev_cap) {
			ev_out[off + k] = __double2float_rn(__ddiv_rn(__dsub_rn((double)raw[k], mean), sd));           // :183
			if (ev_pos) ev_pos[off + k] = (k > 0) ? peak_at(k - 1) : 0u;   // first sample of the event (not part of the reference's output)
		}
}

} // namespace evp

// host side: IEEE fp32 division / fp64 sqrt and division, exactly the operations of revent.c:69 on the clamped variance
static evp::TsWin make_tswin(uint32_t w)
{
	evp::TsWin W;
	W.wf = (float)w; W.rcp = 1.0f / W.wf;
	volatile float vmin = FLT_MIN / W.wf;          // subnormal for w >= 2: x86 SSE keeps it (no flush-to-zero on the host)
	W.vw_min = vmin;
	W.rs_min = 1.0 / sqrt((double)vmin);
	return W;
}

bool events_pipe_supported(uint32_t w1, uint32_t w2) { return w1 >= 1 && w2 >= 1 && w1 <= kPipeMaxWindow && w2 <= kPipeMaxWindow; }

// RA_K1_TRACE: api.cu points this at five timing events (after prefix, t-stat, state machine, scan, finish) per launch
cudaEvent_t *g_pipe_trace = nullptr;

// Segments the state machine cuts a chunk into.  Splitting shortens the recurrence (latency) at the price of ~7 % more
// instructions and four times as many high-priority warps: measured on one box it takes K1 of a 4096-read batch from 2.18
// to 2.04 ms when run alone but lowers the throughput of two batches in flight by 9 % (profiles/r01_ab_fsm_segments.txt).
// So: launches that cannot fill the GPU anyway (< kFsmSplitBelow chunks) are split four ways, large ones are not.
// RA_FSM_SEG = 1 / 2 / 4 forces a count, RA_FSM_FORCE_REPLAY = 1 sends every split chunk through the replay (test hooks).
constexpr uint32_t kFsmSplitBelow = 2048;
static int g_fsm_seg = -1, g_fsm_force_replay = 0;
static int fsm_seg_override()
{
	if (g_fsm_seg < 0) {
		int v = 0;
		if (const char *e = getenv("RA_FSM_SEG")) { const int u = atoi(e); if (u == 1 || u == 2 || u == 4) v = u; }
		if (const char *e = getenv("RA_FSM_FORCE_REPLAY")) g_fsm_force_replay = atoi(e) != 0;
		g_fsm_seg = v;
	}
	return g_fsm_seg;
}
int events_pipe_segments() { return evp::kFsmMaxSeg; }               // what the scratch layout provides for

void events_pipe_layout(uint32_t max_len, const EvParams &p, uint32_t *stride, uint32_t *peak_stride, size_t *floats_per_chunk)
{
	*stride = (max_len + 31u) & ~31u;
	if (*stride == 0) *stride = 32;
	// a detector emits at i >= peak_pos + w/2 + 1 and can only start its next peak at a later sample, so the emissions of one
	// detector are >= w/2 + 2 apart; a segment is at most ceil(max_len / n_seg) rounded up to a tile
	const uint32_t n_seg = (uint32_t)events_pipe_segments();
	const uint32_t seg_len = ((max_len + n_seg - 1) / n_seg + 31u) & ~31u;
	const uint32_t segcap = ((seg_len / (p.w1 / 2 + 2) + seg_len / (p.w2 / 2 + 2) + 4) + 3u) & ~3u;
	*peak_stride = n_seg * segcap;
	// + segment counts; a multiple of 32 floats, so that the scratch of every chunk group starts on a 128-byte line (callers
	// offset by first_chunk * floats_per_chunk) and the 128-byte rows of the stages stay line- and sector-aligned
	*floats_per_chunk = (4 * (size_t)*stride + 2 * (size_t)*peak_stride + evp::kFsmMaxSeg + 31u) & ~(size_t)31u;
}

cudaError_t launch_events_pipe(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len, uint32_t n_chunks,
                               uint32_t max_len, const EvParams &p, int contract, float *scratch, uint32_t *n_ev,
                               uint64_t *ev_off, int chained, float *ev_out, uint64_t ev_cap, uint32_t *ev_pos, cudaStream_t s,
                               cudaEvent_t after_detect, cudaEvent_t scan_wait, cudaEvent_t scan_done, const PipeSplit *split)
{
	using namespace evp;
	if (n_chunks == 0) return cudaSuccess;
	uint32_t stride, pstride; size_t fpc;
	events_pipe_layout(max_len, p, &stride, &pstride, &fpc);
	float *psx = scratch;
	float *pssx = psx + (size_t)n_chunks * stride;
	float *t1 = pssx + (size_t)n_chunks * stride;
	float *t2 = t1 + (size_t)n_chunks * stride;
	uint32_t *peaks = reinterpret_cast<uint32_t *>(t2 + (size_t)n_chunks * stride);
	float *raw = reinterpret_cast<float *>(peaks + (size_t)n_chunks * pstride);
	uint32_t *segcnt = reinterpret_cast<uint32_t *>(raw + (size_t)n_chunks * pstride);
	const int n_seg = fsm_seg_override() ? fsm_seg_override() : (n_chunks < kFsmSplitBelow ? kFsmMaxSeg : 1);
	const uint32_t fsm_ctas = (n_chunks * (uint32_t)n_seg + 32u * kLaneWarps - 1u) / (32u * kLaneWarps);
	const uint32_t lane_ctas = (n_chunks + 32u * kLaneWarps - 1u) / (32u * kLaneWarps);
	constexpr size_t kPrefixSmem = sizeof(PrefixSmem) * kLaneWarps, kFsmSmem = sizeof(FsmSmem) * kLaneWarps;
	static unsigned long long attr_devices = 0;                       // > 48 KB of dynamic shared memory is opt-in, per function and device
	int dev = 0;
	cudaGetDevice(&dev);
	if (!((attr_devices >> (dev & 63)) & 1ull)) {
		cudaError_t ae = cudaFuncSetAttribute(k_prefix_pipe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPrefixSmem);
		if (ae == cudaSuccess) ae = cudaFuncSetAttribute(k_prefix_pipe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPrefixSmem);
		if (ae == cudaSuccess) ae = cudaFuncSetAttribute(k_fsm_pipe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFsmSmem);
		if (ae == cudaSuccess) ae = cudaFuncSetAttribute(k_fsm_pipe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFsmSmem);
		if (ae == cudaSuccess) ae = cudaFuncSetAttribute(k_fsm_pipe<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFsmSmem);
		if (ae != cudaSuccess) return ae;
		__atomic_fetch_or(&attr_devices, 1ull << (dev & 63), __ATOMIC_RELAXED);   // idempotent, so a race between host threads is harmless
	}
	const dim3 tgrid((max_len + kTsBlock - 1) / kTsBlock, n_chunks);
	const TsWin W1 = make_tswin(p.w1), W2 = make_tswin(p.w2);
	// The two lane-per-chunk kernels are latency-bound and small (one CTA per 128 chunks); with `split` they go to a
	// high-priority stream so that their CTAs are placed ahead of the pending CTAs of other groups' t-stat grids.
	cudaStream_t sl = split ? split->hi : s;
	if (split) { cudaEventRecord(split->e[0], s); cudaStreamWaitEvent(sl, split->e[0], 0); }
	if (contract) k_prefix_pipe<true><<<lane_ctas, 32 * kLaneWarps, kPrefixSmem, sl>>>(sig, chunk_off, chunk_len, n_chunks, psx, pssx, stride);
	else k_prefix_pipe<false><<<lane_ctas, 32 * kLaneWarps, kPrefixSmem, sl>>>(sig, chunk_off, chunk_len, n_chunks, psx, pssx, stride);
	if (g_pipe_trace) cudaEventRecord(g_pipe_trace[0], sl);
	if (split) { cudaEventRecord(split->e[1], sl); cudaStreamWaitEvent(s, split->e[1], 0); }
	if (contract) k_tstat_pipe<true><<<tgrid, kTsThreads, 0, s>>>(chunk_len, psx, pssx, stride, t1, t2, p.w1, p.w2, W1, W2);
	else k_tstat_pipe<false><<<tgrid, kTsThreads, 0, s>>>(chunk_len, psx, pssx, stride, t1, t2, p.w1, p.w2, W1, W2);
	if (g_pipe_trace) cudaEventRecord(g_pipe_trace[1], s);
	if (split) { cudaEventRecord(split->e[2], s); cudaStreamWaitEvent(sl, split->e[2], 0); }
	if (n_seg == 4) k_fsm_pipe<4><<<fsm_ctas, 32 * kLaneWarps, kFsmSmem, sl>>>(chunk_len, n_chunks, t1, t2, stride, p, peaks, pstride, n_ev, segcnt, g_fsm_force_replay);
	else if (n_seg == 2) k_fsm_pipe<2><<<fsm_ctas, 32 * kLaneWarps, kFsmSmem, sl>>>(chunk_len, n_chunks, t1, t2, stride, p, peaks, pstride, n_ev, segcnt, g_fsm_force_replay);
	else k_fsm_pipe<1><<<fsm_ctas, 32 * kLaneWarps, kFsmSmem, sl>>>(chunk_len, n_chunks, t1, t2, stride, p, peaks, pstride, n_ev, segcnt, 0);
	if (g_pipe_trace) cudaEventRecord(g_pipe_trace[2], sl);
	if (split) { cudaEventRecord(split->e[3], sl); cudaStreamWaitEvent(s, split->e[3], 0); }
	if (after_detect) cudaEventRecord(after_detect, s);
	// CSR offsets continue from the previous launch group: when that group runs on another stream, only the scan waits for it.
	// (Traced with RA_K1_TRACE: in a large batch run alone the one-CTA scans and the finish kernels sit behind the pending
	// t-stat grids of the other groups; moving the scan to the high-priority stream frees the scans but not the finish
	// kernels, and costs 3 % of the throughput of two batches in flight -- profiles/r01_ab_fsm_segments.txt.)
	if (scan_wait) cudaStreamWaitEvent(s, scan_wait, 0);
	cudaError_t e = launch_exclusive_scan_u32_chained(n_ev, ev_off, n_chunks, chained, s);
	if (e != cudaSuccess) return e;
	if (scan_done) cudaEventRecord(scan_done, s);
	if (g_pipe_trace) cudaEventRecord(g_pipe_trace[3], s);
	const unsigned fblocks = (unsigned)(((uint64_t)n_chunks * 32 + 127) / 128);
#define RA_FINISH(C, SG) k_finish_pipe<C, SG><<<fblocks, 128, 0, s>>>(chunk_len, n_chunks, psx, stride, peaks, pstride, n_ev, ev_off, ev_out, ev_cap, raw, ev_pos, segcnt, n_seg)
	if (contract) { if (n_seg > 1) RA_FINISH(true, true); else RA_FINISH(true, false); }
	else { if (n_seg > 1) RA_FINISH(false, true); else RA_FINISH(false, false); }
#undef RA_FINISH
	if (g_pipe_trace) cudaEventRecord(g_pipe_trace[4], s);
	return cudaGetLastError();
}

} // namespace ra
