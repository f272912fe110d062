// dtw_core.cuh -- device-side DTW building blocks shared by the item kernels (dtw.cu) and the chain
// kernel (chains.cu).
//
// Semantics (bit-identical to the reference for every schedule, because each cell is
// min(min(top,left),topleft) + |a-b| -- min is exact, one rounding per cell):
//   banded : the cell set of DTW_global_slantedbanded_antidiagonalwise, src/dtw.cpp:273-520
//            (closed form in oracle/ra_oracle.c:ora_dtw_banded and DESIGN.md), out-of-set = 1e10f;
//   full   : DTW_global, src/dtw.cpp:37-66 (whole rectangle).
#pragma once
#include "ra_common.cuh"

namespace ra {

constexpr float kInf = 1e10f;

// Enumerates the anti-diagonals of one problem in order.  n >= m for banded problems (the caller
// swaps, dtw.cpp:284-292); rows j index the shorter sequence.
struct DiagIter {
	int n, m, P, S;
	bool full;
	int c, r, d;
	bool pending_primary;
	int lo, hi;       // rows of the current diagonal, clipped to the matrix; lo > hi = empty

	__device__ __forceinline__ void set_range(int start_i, int start_j, int len)
	{
		// cells (start_i - o, start_j + o), o in [o0, o1)   dtw.cpp:365-366, :419-420
		const int o0 = max(max(0, start_i - (n - 1)), -start_j);
		const int o1 = min(min(len, start_i + 1), m - start_j);
		lo = start_j + o0; hi = start_j + o1 - 1;
		d = start_i + start_j;
	}
	__device__ __forceinline__ void init(int n_, int m_, int band_radius)
	{
		n = n_; m = m_;
		full = band_radius < 0;
		if (!full) {
			// dtw.cpp:298-302 (32-bit unsigned arithmetic in the reference)
			const uint32_t extra = (((uint32_t)n - (uint32_t)m) * (uint32_t)band_radius + (uint32_t)n - 1u) / (uint32_t)n;
			const int R = band_radius + (int)extra;
			P = R + ((R % 2 == 0) ? 1 : 0);
			S = R + ((R % 2 == 1) ? 1 : 0);
		} else { P = S = 0; }
		c = 0; r = 0; d = 0; lo = 0; hi = 0; pending_primary = false;   // iteration 0: the corner cell only (:317-347)
	}
	__device__ __forceinline__ bool next()
	{
		if (full) {
			if (++d > n + m - 2) return false;
			lo = max(0, d - (n - 1)); hi = min(m - 1, d);
			return true;
		}
		if (pending_primary) { pending_primary = false; set_range(c + P / 2, r - P / 2, P); return true; }
		if (++c >= n) return false;
		if ((int64_t)(r + 1) * (int64_t)n <= (int64_t)m * (int64_t)c) {   // :352-359
			++r; pending_primary = true;
			set_range(c + S / 2 - 1, r - S / 2, S);                        // secondary, :362-366
			return true;
		}
		set_range(c + P / 2, r - P / 2, P);                                // primary, :416-420
		return true;
	}
	__device__ __forceinline__ int max_width() const { return full ? m : max(P, S); }
};

__device__ __forceinline__ int band_width_of(int n, int m, int band_radius)
{
	if (band_radius < 0) return m;
	const uint32_t extra = (((uint32_t)n - (uint32_t)m) * (uint32_t)band_radius + (uint32_t)n - 1u) / (uint32_t)n;
	return band_radius + (int)extra + 1;
}

// Warp-cooperative cost of one problem: lanes stride over the rows of each anti-diagonal; three rolling
// diagonals in `ring` (3*cap floats, cap a power of two >= the widest diagonal; shared or global memory).
// A is the longer sequence (n >= m) for banded problems.  Returns dp(n-1, m-1) in every lane and adds the
// number of evaluated cells to *cells (lane-uniform).
__device__ __forceinline__ float warp_dtw_cost(const float *__restrict__ A, int n, const float *__restrict__ B, int m,
                                               int band, float *ring, uint32_t cap, int lane, uint64_t *cells)
{
	DiagIter di; di.init(n, m, band);
	const uint32_t mask = cap - 1;
	float *r0 = ring, *r1 = ring + cap, *r2 = ring + 2 * (size_t)cap;
	int lo1 = 0, hi1 = -1, lo2 = 0, hi2 = -1;
	if (lane == 0) r0[0] = fabsf(__ldg(A) - __ldg(B));
	__syncwarp();
	int lo0 = 0, hi0 = 0;
	uint64_t cnt = 1;
	for (;;) {
		float *t = r2; r2 = r1; r1 = r0; r0 = t;
		lo2 = lo1; hi2 = hi1; lo1 = lo0; hi1 = hi0;
		if (!di.next()) break;
		lo0 = di.lo; hi0 = di.hi;
		const int d = di.d;
		if (hi0 >= lo0) cnt += (uint64_t)(hi0 - lo0 + 1);
		for (int j = lo0 + lane; j <= hi0; j += 32) {
			const int i = d - j;
			// (i, j-1) and (i-1, j) on diagonal d-1; (i-1, j-1) on diagonal d-2
			const float top = (j - 1 >= lo1 && j - 1 <= hi1) ? r1[(uint32_t)(j - 1) & mask] : kInf;
			const float left = (j >= lo1 && j <= hi1) ? r1[(uint32_t)j & mask] : kInf;
			const float tl = (j - 1 >= lo2 && j - 1 <= hi2) ? r2[(uint32_t)(j - 1) & mask] : kInf;
			r0[(uint32_t)j & mask] = fminf(fminf(top, left), tl) + fabsf(__ldg(A + i) - __ldg(B + j));
		}
		__syncwarp();
	}
	if (cells) *cells += cnt;
	const float res = r1[(uint32_t)(m - 1) & mask];      // after the final rotation the last diagonal is r1
	__syncwarp();
	return res;
}

// ---------------------------------------------------------------------------------------------------
// Row form of the banded cell set.  For A = longer sequence (index i, n entries), B = shorter
// (index j, m entries), R' = slant-extended radius (dtw.cpp:298-300), odd = R' & 1,
// h = odd ? (R'-1)/2 : R'/2, the cell set of dtw.cpp:273-520 is, row by row,
//     jhi(i) = (i+h+odd > n-1) ? m-1 : min(m-1, floor(m*(i+h+odd)/n) + h)
//     jlo(i) = (i-h-odd < 0)   ? 0   : max(0,   floor(m*(i-h-odd)/n) - h)
// contiguous in every row and non-decreasing in i (tests/test_band_rows.py checks this closed form
// against the anti-diagonal enumeration; tools/band_rows_check.py is the same check as a script).
// Full matrix: jlo = 0, jhi = m-1.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int slant_radius(int n, int m, int band_radius)
{
	// dtw.cpp:298-300: R' = R + ceil((n-m)*R / n), 32-bit unsigned arithmetic.  ceil(x/n) = #{k >= 0 : k*n < x} and
	// x < R*n, so for the radii that dominate (R <= 4) four compares replace the division.
	const uint32_t x = ((uint32_t)n - (uint32_t)m) * (uint32_t)band_radius, un = (uint32_t)n;
	if (band_radius <= 4 && un < (1u << 28)) return band_radius + (int)(x > 0u) + (int)(x > un) + (int)(x > 2u * un) + (int)(x > 3u * un);
	return band_radius + (int)((x + un - 1u) / un);
}

// Thread-private cost of one NARROW problem: one row of A per step, the previous row and the B values
// of the live columns in two lane-interleaved shared-memory rings of kNarrowRing entries
// (ring[slot*32 + lane], bank-conflict free).  Requires 2*R'+4 <= kNarrowRing (band) or m+2 <= kNarrowRing (full).
constexpr int kNarrowRing = 16;
__device__ __forceinline__ bool narrow_ok(int n, int m, int band_radius)
{
	if (band_radius < 0) return m + 2 <= kNarrowRing;
	return 2 * slant_radius(n, m, band_radius) + 4 <= kNarrowRing;
}

struct LdgLoader {          // operands straight from global memory (read-only path)
	const float *__restrict__ p;
	__device__ __forceinline__ float operator[](int k) const { return __ldg(p + k); }
};
struct SmemLoader {         // operands staged in shared memory by the CTA
	const float *p;
	__device__ __forceinline__ float operator[](int k) const { return p[k]; }
};

template <class LA, class LB>
__device__ __forceinline__ float thread_dtw_narrow(const LA A, int n, const LB B, int m,
                                                   int band_radius, float *row /* + lane */, float *bv /* + lane */, uint32_t *cells)
{
	constexpr uint32_t M = kNarrowRing - 1;
	int h = 0, lead = 0;                 // lead = h + odd: how far the column trackers run ahead of / behind the row
	const bool full = band_radius < 0;
	if (!full) {
		const int Rp = slant_radius(n, m, band_radius);
		const int odd = Rp & 1;
		h = odd ? (Rp - 1) / 2 : Rp / 2;
		lead = h + odd;
	}
	// upper tracker: x_hi = i + lead, r_hi = floor(m*x_hi/n), e_hi = (m*x_hi) mod n
	int r_hi = 0, e_hi = 0;
	for (int s = 0; s < lead; ++s) { e_hi += m; if (e_hi >= n) { e_hi -= n; ++r_hi; } }
	int r_lo = 0, e_lo = 0;              // lower tracker, valid once i - lead >= 0
	int pjh = -1;                        // jhi of the previous row (columns <= pjh hold dp[i-1][.])
	int pjl = 0;
	float last = kInf;
	uint32_t cnt = 0;
	for (int i = 0; i < n; ++i) {
		int jh = m - 1, jl = 0;
		if (!full) {
			if (i + lead <= n - 1) jh = min(m - 1, r_hi + h);
			if (i - lead >= 0) jl = max(0, r_lo - h);
		}
		// columns entering the band: fetch B, no value from the previous row
		for (int j = pjh + 1; j <= jh; ++j) { bv[((uint32_t)j & M) * 32] = B[j]; row[((uint32_t)j & M) * 32] = kInf; }
		const float a_i = A[i];
		// dp(i-1, jl-1): inside the previous row's range or outside the set; row 0 starts from 0 so that dp(0,0) = |A0-B0|
		float diag = (i == 0) ? 0.0f : ((jl - 1 >= pjl) ? row[((uint32_t)(jl - 1) & M) * 32] : kInf);
		float left = kInf;
		for (int j = jl; j <= jh; ++j) {
			const uint32_t s = ((uint32_t)j & M) * 32;
			const float up = row[s];
			const float v = fminf(fminf(left, up), diag) + fabsf(a_i - bv[s]);
			row[s] = v; diag = up; left = v;
		}
		cnt += (uint32_t)(jh - jl + 1);
		last = left;
		pjh = jh; pjl = jl;
		if (!full) {
			e_hi += m; if (e_hi >= n) { e_hi -= n; ++r_hi; }
			if (i - lead >= 0) { e_lo += m; if (e_lo >= n) { e_lo -= n; ++r_lo; } }
		}
	}
	if (cells) *cells += cnt;
	return last;                         // dp(n-1, m-1): jhi(n-1) = m-1 always
}

// ---------------------------------------------------------------------------------------------------
// Register-resident anti-diagonal evaluator for a compile-time slant-extended radius RP (= R').
//
// This is the reference's own schedule (dtw.cpp:352-485): per column c of the longer sequence one
// primary anti-diagonal P(c) = {(c+k, r-k), |k| <= h} (2h+1 cells) and, when r = floor(m*c/n) has just
// incremented, a secondary S(c) before it.  Every diagonal depends on the two before it with FIXED
// index offsets (derived in DESIGN.md "K2 register evaluator"), so the three live diagonals, the
// A window [c-h-1, c+h] and the B window [r-h-1, r+h] are plain registers and the code of one column is
// straight-line: no loop over cells, no shared-memory round trip, no per-cell bounds test.
// Cells outside the matrix are neutralised by operand sentinels (A = +1e18 before/after the sequence,
// B = -1e18), which makes their cost, hence their value, >= 1e18; cells outside the band are the
// constant kInf operands below.  Because min() discards them wherever a real predecessor exists and
// the centre staircase (c, r(c)) is always connected, dp(n-1, m-1) is bit-identical to the reference's
// (whose out-of-band value is 1e10f) -- asserted against the oracle in tests/test_gpu_parity.py.
// ---------------------------------------------------------------------------------------------------
//
// MIXED (RP == 2 only): lanes whose problem has R' = 1 run the same code with `r1` set.  The R' = 1 cell set is the R' = 2
// set without the two outer cells of every primary diagonal (same secondaries, same centre cell), so forcing those two
// values to kInf -- what the reference holds for an out-of-set cell -- evaluates the R' = 1 problem exactly.
#ifndef RA_DIAG_UNROLL
#define RA_DIAG_UNROLL 2
#endif
constexpr int kDiagUnroll = RA_DIAG_UNROLL;
template <int RP, bool COUNT, bool MIXED = false, class LA, class LB>
__device__ __forceinline__ float thread_dtw_diag(const LA A, int n, const LB B, int m, uint32_t *cells, bool r1 = false)
{
	static_assert(!MIXED || RP == 2, "MIXED is the R' = 1 inside R' = 2 embedding");
	static_assert(!(MIXED && COUNT), "cell counting of mixed rounds is done outside the evaluator (diag_cells, chains_sparse.cu)");
	constexpr bool ODD = (RP & 1) != 0;
	constexpr int h = ODD ? (RP - 1) / 2 : RP / 2;
	constexpr int NP = 2 * h + 1;                 // primary length (odd)
	constexpr int NS = ODD ? 2 * h + 2 : 2 * h;   // secondary length (even)
	constexpr int NW = 2 * h + 2;                 // operand windows and the d-2 diagonal
	constexpr float SA = 1e18f, SB = -1e18f;
	float pold[NP], xq[NW], aw[NW], bw[NW];
#pragma unroll
	for (int t = 0; t < NP; ++t) pold[t] = kInf;
#pragma unroll
	for (int u = 0; u < NW; ++u) {
		xq[u] = (u == h + 1) ? 0.0f : kInf;       // virtual dp(-1,-1) = 0 so that dp(0,0) = |A0 - B0|
		const int i = -h - 1 + u;                 // aw[u] = A[c-h-1+u], bw[u] = B[r-h-1+u] at c = 0, r = 0
		aw[u] = (i >= 0 && i < n) ? A[i] : SA;
		bw[u] = (i >= 0 && i < m) ? B[i] : SB;
	}
	int r = 0, e = 0;
	uint32_t cnt = 0;
#pragma unroll kDiagUnroll
	for (int c = 0; c < n; ++c) {
		bool inc = false;
		if (c > 0) {                                                  // same for every lane that is still running
			e += m; inc = e >= n; e -= inc ? n : 0; r += inc ? 1 : 0;  // dtw.cpp:352-359
#pragma unroll
			for (int u = 0; u < NW - 1; ++u) aw[u] = aw[u + 1];
			aw[NW - 1] = (c + h < n) ? A[c + h] : SA;
			// lanes differ in inc at every column: the B window moves by selects, not by a branch
			const float bnew = (r + h < m) ? B[min(r + h, m - 1)] : SB;
#pragma unroll
			for (int u = 0; u < NW - 1; ++u) bw[u] = inc ? bw[u + 1] : bw[u];
			bw[NW - 1] = inc ? bnew : bw[NW - 1];
		}
		// ---- secondary diagonal S(c), only meaningful when inc ----
		float s[NS > 0 ? NS : 1];
#pragma unroll
		for (int t = 0; t < NS; ++t) {
			float pa, pb, tl, cost;
			if (ODD) {        // cell i = c-h-1+t, j = r+h-t
				pa = (t >= 1) ? pold[t - 1] : kInf;
				pb = (t <= 2 * h) ? pold[t] : kInf;
				tl = xq[t];
				cost = fabsf(aw[t] - bw[2 * h + 1 - t]);
			} else {          // cell i = c-h+t, j = r-1+h-t
				pa = pold[t]; pb = pold[t + 1];
				tl = xq[t + 1];
				cost = fabsf(aw[t + 1] - bw[2 * h - t]);
			}
			s[t] = fminf(fminf(pa, pb), tl) + cost;
		}
		// ---- primary diagonal P(c): cell i = c-h+t, j = r+h-t ----
		float pnew[NP];
#pragma unroll
		for (int t = 0; t < NP; ++t) {
			// after S(c): prev = S(c), prev2 = P(c-1); otherwise prev = P(c-1), prev2 = xq
			float sa, sb;
			if (ODD) { sa = s[t]; sb = s[t + 1]; }
			else { sa = (t >= 1) ? s[t - 1] : kInf; sb = (t <= 2 * h - 1) ? s[t] : kInf; }
			const float na = pold[t], nb = (t + 1 <= 2 * h) ? pold[t + 1] : kInf, ntl = xq[t + 1];
			const float pa = inc ? sa : na, pb = inc ? sb : nb, tl = inc ? pold[t] : ntl;
			pnew[t] = fminf(fminf(pa, pb), tl) + fabsf(aw[t + 1] - bw[2 * h + 1 - t]);
			if (MIXED && t != h) pnew[t] = r1 ? kInf : pnew[t];
		}
		// ---- the d-2 diagonal of the next column, anchored at i = c-h-1 ----
#pragma unroll
		for (int u = 0; u < NW; ++u) {
			float fromS;
			if (ODD) fromS = s[u];
			else fromS = (u >= 1 && u <= 2 * h) ? s[u - 1] : kInf;
			const float fromP = (u <= 2 * h) ? pold[u] : kInf;
			xq[u] = inc ? fromS : fromP;
		}
#pragma unroll
		for (int t = 0; t < NP; ++t) pold[t] = pnew[t];
		if (COUNT) {
			// cells of P(c) and S(c) inside the matrix (dtw.cpp:365-366, :419-420)
			const int klo = max(max(-h, -c), r - (m - 1)), khi = min(min(h, n - 1 - c), r);
			cnt += (uint32_t)max(0, khi - klo + 1);
			if (inc) {
				// S(c): (c+k, r-1-k), k in [-NS/2, NS/2-1]
				const int slo = max(max(-(NS / 2), -c), r - 1 - (m - 1)), shi = min(min(NS / 2 - 1, n - 1 - c), r - 1);
				cnt += (uint32_t)max(0, shi - slo + 1);
			}
		}
	}
	if (COUNT && cells) *cells += cnt;
	return pold[h];                               // dp(n-1, m-1) = P(n-1), k = 0
}

// Lane-private cost of one SMALL problem (widest diagonal <= kLaneRing): the three rolling diagonals
// live in shared memory, interleaved across the warp ([slot][lane], bank-conflict free).
constexpr int kLaneRing = 8;
__device__ __forceinline__ float lane_dtw_cost(const float *__restrict__ A, int n, const float *__restrict__ B, int m,
                                               int band, float *ring_lane /* + slot*32 */, uint32_t *cells)
{
	DiagIter di; di.init(n, m, band);
	constexpr uint32_t mask = kLaneRing - 1;
	int o0 = 0, o1 = kLaneRing, o2 = 2 * kLaneRing;      // slot offsets of the current / previous / before-previous diagonal
	int lo1 = 0, hi1 = -1, lo2 = 0, hi2 = -1, lo0 = 0, hi0 = 0;
	ring_lane[(o0 + 0) * 32] = fabsf(__ldg(A) - __ldg(B));
	uint32_t cnt = 1;
	for (;;) {
		const int t = o2; o2 = o1; o1 = o0; o0 = t;
		lo2 = lo1; hi2 = hi1; lo1 = lo0; hi1 = hi0;
		if (!di.next()) break;
		lo0 = di.lo; hi0 = di.hi;
		const int d = di.d;
		for (int j = lo0; j <= hi0; ++j) {
			const int i = d - j;
			const float top = (j - 1 >= lo1 && j - 1 <= hi1) ? ring_lane[(o1 + ((uint32_t)(j - 1) & mask)) * 32] : kInf;
			const float left = (j >= lo1 && j <= hi1) ? ring_lane[(o1 + ((uint32_t)j & mask)) * 32] : kInf;
			const float tl = (j - 1 >= lo2 && j - 1 <= hi2) ? ring_lane[(o2 + ((uint32_t)(j - 1) & mask)) * 32] : kInf;
			ring_lane[(o0 + ((uint32_t)j & mask)) * 32] = fminf(fminf(top, left), tl) + fabsf(__ldg(A + i) - __ldg(B + j));
			++cnt;
		}
	}
	if (cells) *cells += cnt;
	return ring_lane[(o1 + ((uint32_t)(m - 1) & mask)) * 32];
}

} // namespace ra
