// dtw_wave.cu -- K2/K3 for medium and large problems: one warp per DTW problem, register-resident skewed wavefront.
//
// Same semantics as dtw.cu (bit-identical to the reference for every schedule: each cell is
// min(min(top,left),topleft) + |a-b|, min is exact, one rounding per cell):
//   banded : the cell set of DTW_global_slantedbanded_antidiagonalwise, src/dtw.cpp:273-520, out-of-set = 1e10f
//   full   : DTW_global, src/dtw.cpp:37-66
//
// Layout.  A = the longer sequence (index i, n entries, "columns"), B = the shorter (index j, m entries, "rows").  In row
// form (dtw_core.cuh) the band of column i is the contiguous row range [jlo(i), jhi(i)], both ends non-decreasing in i, at
// most W = 2R'+2 rows tall.  Rows are cut into groups of H consecutive rows; lane t owns the groups g = t, t+32, t+64, ...
// (cyclic), keeping the group's H reference values b[j] and the H cells of its previous column in REGISTERS.  At step s
// group g works on column i = s - g (skew of one column per group), so the cell above a group's first row (computed by
// the neighbouring lane one step earlier) and the one diagonally above (two steps earlier) arrive by ONE rotate-shuffle
// per step -- one shuffle per H cells, against one shared-memory round trip per cell in the generic kernel.  With
// W <= 31*H a lane's consecutive groups are never active at the same step (proof in DESIGN.md), so a single warp sweeps
// the whole band at ~W/(32H) lane efficiency.  Per cell: |a-b| (1), min3 (2), add (1), band mask (2-3).
//
// Compiled with -fmad=false (nothing here may be contracted; there is no multiply on the path anyway).
#include "dtw_core.cuh"

namespace ra {

// jlo(i), jhi(i) of the row form of the banded cell set (dtw_core.cuh; tests/test_band_rows.py), advanced column by column
struct BandTrack {
	int n, m, h, lead;        // lead = h + odd
	bool full;
	int r_hi, e_hi, r_lo, e_lo, i;
	__device__ __forceinline__ void setup(int n_, int m_, int band)
	{
		n = n_; m = m_; full = band < 0; h = 0; lead = 0;
		if (!full) {
			const int Rp = slant_radius(n, m, band);
			const int odd = Rp & 1;
			h = odd ? (Rp - 1) / 2 : Rp / 2;
			lead = h + odd;
		}
		r_hi = e_hi = r_lo = e_lo = 0; i = 0;
	}
	__device__ __forceinline__ void init(int i_)
	{
		i = i_;
		if (full) return;
		const long long xh = (long long)i + lead, xl = (long long)i - lead;
		r_hi = (int)((long long)m * xh / n); e_hi = (int)((long long)m * xh % n);
		if (xl > 0) { r_lo = (int)((long long)m * xl / n); e_lo = (int)((long long)m * xl % n); }
		else { r_lo = 0; e_lo = 0; }
	}
	__device__ __forceinline__ void step()       // column i -> i + 1
	{
		++i;
		if (full) return;
		e_hi += m; if (e_hi >= n) { e_hi -= n; ++r_hi; }
		if (i - lead > 0) { e_lo += m; if (e_lo >= n) { e_lo -= n; ++r_lo; } }
	}
	__device__ __forceinline__ int jhi() const { return (full || i + lead > n - 1) ? m - 1 : min(m - 1, r_hi + h); }
	__device__ __forceinline__ int jlo() const { return (full || i - lead < 0) ? 0 : max(0, r_lo - h); }
};

// One problem swept by T = 32*WARPS threads (thread index t).  The value a group's first row needs from the row above it
// travels lane to lane by shuffle; between warps (WARPS > 1) through a double-buffered shared-memory mailbox, one barrier per
// step.  Returns dp(n-1, m-1) in the thread that owns it (*owner = true there).
// FULL (the whole matrix, DTW_global): no band to track and no cell to mask -- rows j >= m of the last group hold garbage that
// only flows downwards (a row reads the row above it, never the one below) and is never read by dp(n-1, m-1).
template <int H, int WARPS, bool FULL = false>
__device__ __forceinline__ float wave_cost(const float *__restrict__ A, int n, const float *__restrict__ B, int m, int band, int t,
                                           float (*mbox)[WARPS], bool *owner)
{
	constexpr int T = 32 * WARPS;
	const int lane = t & 31, warp = t >> 5;
	BandTrack tr; tr.setup(n, m, band);
	int g = t;                          // current row group of this thread
	int i = -t;                         // its column at step s: i = s - g
	bool need_init = true;
	float prev[H], bv[H];
#pragma unroll
	for (int r = 0; r < H; ++r) { prev[r] = kInf; const int j = g * H + r; bv[r] = (j < m) ? __ldg(B + j) : 0.0f; }
	float recv = kInf, recv_prev = kInf, result = kInf;
	int jlo = 0, jhi = -1, jlo_p = 0, jhi_p = -1;          // bounds of this thread's current and previous column
	const int g_final = (m - 1) / H, r_final = (m - 1) % H;
	const int steps = n + g_final;
	*owner = false;
	for (int s = 0; s < steps; ++s, ++i) {
		bool act = false;
#pragma unroll 1
		for (int attempt = 0; attempt < 2; ++attempt) {
			if (g * H >= m || i < 0) break;                  // no group left, or this group has not reached column 0 yet
			bool finished = (i >= n);
			if (FULL) {
				if (!finished) { act = true; break; }
			} else if (!finished) {
				if (need_init) {
					if (i > 0) { tr.init(i - 1); jlo_p = tr.jlo(); jhi_p = tr.jhi(); tr.step(); }
					else { tr.init(0); jlo_p = 0; jhi_p = -1; }
					need_init = false;
				} else if (attempt == 0) { jlo_p = jlo; jhi_p = jhi; tr.step(); }
				jlo = tr.jlo(); jhi = tr.jhi();
				finished = jlo > g * H + H - 1;              // the band has moved past this group's rows
			}
			if (!finished) { act = jhi >= g * H; break; }
			// next group of this thread (cyclic ownership); it may become active at this very step
			g += T; i -= T; need_init = true;
#pragma unroll
			for (int r = 0; r < H; ++r) { prev[r] = kInf; const int j = g * H + r; bv[r] = (j < m) ? __ldg(B + j) : 0.0f; }
		}
		float send = kInf;
		if (act) {
			const int jt = g * H - 1;                        // the row above this group's first row (owned by the neighbouring thread)
			float up = (g > 0 && (FULL || (jt >= jlo && jt <= jhi))) ? recv : kInf;                     // dp(i, jt)
			float diag = (g > 0 && i > 0 && (FULL || (jt >= jlo_p && jt <= jhi_p))) ? recv_prev : kInf;   // dp(i-1, jt)
			if (g == 0 && i == 0) diag = 0.0f;               // virtual dp(-1,-1) = 0 so that dp(0,0) = |A0 - B0|
			const float a_i = __ldg(A + i);
			const unsigned base = (unsigned)(g * H - jlo), wdt = (unsigned)(jhi - jlo);
#pragma unroll
			for (int r = 0; r < H; ++r) {
				const float left = prev[r];
				float v = fminf(fminf(up, left), diag) + fabsf(a_i - bv[r]);
				if (!FULL) v = (base + (unsigned)r <= wdt) ? v : kInf;   // rows outside [jlo, jhi] are outside the cell set
				diag = left; up = v; prev[r] = v;
			}
			send = up;
			if (g == g_final && i == n - 1) {                // dp(n-1, m-1)
#pragma unroll
				for (int r = 0; r < H; ++r) if (r == r_final) result = prev[r];
				*owner = true;
			}
		}
		recv_prev = recv;
		if (WARPS == 1) recv = __shfl_sync(0xffffffffu, send, (lane + 31) & 31);
		else {
			float x = __shfl_up_sync(0xffffffffu, send, 1);
			if (lane == 31) mbox[s & 1][warp] = send;
			__syncthreads();
			if (lane == 0) x = mbox[s & 1][(warp + WARPS - 1) % WARPS];
			recv = x;
		}
	}
	return result;
}

constexpr int kWaveWarps = 4;

// one warp per problem, kWaveWarps problems in flight per CTA
template <int H>
__global__ void __launch_bounds__(kWaveWarps * 32)
k_dtw_wave(DtwLaunch L, const uint32_t *__restrict__ order, uint32_t n_order, uint32_t *work_counter)
{
	const int lane = threadIdx.x & 31;
	for (;;) {
		uint32_t slot = 0;
		if (lane == 0) slot = atomicAdd(work_counter, 1u);
		slot = __shfl_sync(0xffffffffu, slot, 0);
		if (slot >= n_order) break;
		const uint32_t k = order[slot];
		const ra_dtw_item_t it = L.items[k];
		const float *A = L.a_arena + it.a_off, *B = L.b_arena + it.b_off;
		int n = (int)it.a_len, m = (int)it.b_len;
		if (n < m) { const float *t = A; A = B; B = t; const int tl = n; n = m; m = tl; }      // dtw.cpp:284-292; a full matrix is symmetric
		bool owner;
		float res = it.band_radius < 0 ? wave_cost<H, 1, true>(A, n, B, m, -1, lane, nullptr, &owner)
		                               : wave_cost<H, 1, false>(A, n, B, m, it.band_radius, lane, nullptr, &owner);
		if (owner) {
			if (it.exclude_last) res -= fabsf(__ldg(A + n - 1) - __ldg(B + m - 1));               // dtw.cpp:514-516 / :60-62
			L.cost[k] = res;
		}
		__syncwarp();
	}
}

// one CTA of kWideWarps warps per problem: bands (or full matrices) up to 255*H rows tall
constexpr int kWideWarps = 8;
template <int H>
__global__ void __launch_bounds__(kWideWarps * 32)
k_dtw_wave_cta(DtwLaunch L, const uint32_t *__restrict__ order, uint32_t n_order, uint32_t *work_counter)
{
	__shared__ float mbox[2][kWideWarps];
	__shared__ uint32_t s_slot;
	for (;;) {
		__syncthreads();                                     // the previous problem's mailbox and slot are no longer read
		if (threadIdx.x == 0) s_slot = atomicAdd(work_counter, 1u);
		__syncthreads();
		const uint32_t slot = s_slot;
		if (slot >= n_order) break;
		const uint32_t k = order[slot];
		const ra_dtw_item_t it = L.items[k];
		const float *A = L.a_arena + it.a_off, *B = L.b_arena + it.b_off;
		int n = (int)it.a_len, m = (int)it.b_len;
		if (n < m) { const float *t = A; A = B; B = t; const int tl = n; n = m; m = tl; }
		bool owner;
		float res = it.band_radius < 0 ? wave_cost<H, kWideWarps, true>(A, n, B, m, -1, (int)threadIdx.x, mbox, &owner)
		                               : wave_cost<H, kWideWarps, false>(A, n, B, m, it.band_radius, (int)threadIdx.x, mbox, &owner);
		if (owner) {
			if (it.exclude_last) res -= fabsf(__ldg(A + n - 1) - __ldg(B + m - 1));
			L.cost[k] = res;
		}
	}
}

// Rows the band can span in one column (W): 2R'+2 for a band, m for the full matrix.  Class = rows per thread H (a power of
// two) with W <= (T-1)*H: 1..32 for one warp per problem (T = 32), 64 + H (H = 8, 16, 32) for one CTA of 256 threads per
// problem; 0 when the problem is taller than that (the generic kernel with its HBM ring takes it).
int dtw_wave_class(uint32_t a_len, uint32_t b_len, int32_t band_radius)
{
	const uint64_t n = a_len > b_len ? a_len : b_len, m = a_len > b_len ? b_len : a_len;
	if (n == 0 || m == 0 || n >= (1u << 30)) return 0;
	uint64_t W = m;
	if (band_radius >= 0) {
		const uint64_t Rp = (uint64_t)band_radius + ((n - m) * (uint64_t)band_radius + n - 1) / n;
		W = 2 * Rp + 2 < m ? 2 * Rp + 2 : m;
	}
	for (int H = 1; H <= 32; H <<= 1)
		if (W <= 31u * (uint64_t)H) return H;
	for (int H = 8; H <= 32; H <<= 1)
		if (W <= (uint64_t)(32 * kWideWarps - 1) * (uint64_t)H) return 64 + H;
	return 0;
}

cudaError_t launch_dtw_wave(const DtwLaunch &L, int cls, const uint32_t *order_dev, uint32_t n_order, uint32_t *work_counter, int n_sm, cudaStream_t s)
{
	if (n_order == 0) return cudaSuccess;
	cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	const uint32_t want = (n_order + kWaveWarps - 1) / kWaveWarps, cap = (uint32_t)n_sm * 12u;
	const uint32_t grid = want < cap ? want : cap;
	const uint32_t wgrid = n_order < (uint32_t)n_sm * 4u ? n_order : (uint32_t)n_sm * 4u;
	switch (cls) {
	case 1: k_dtw_wave<1><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 2: k_dtw_wave<2><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 4: k_dtw_wave<4><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 8: k_dtw_wave<8><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 16: k_dtw_wave<16><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 32: k_dtw_wave<32><<<grid, kWaveWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 64 + 8: k_dtw_wave_cta<8><<<wgrid, kWideWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 64 + 16: k_dtw_wave_cta<16><<<wgrid, kWideWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	case 64 + 32: k_dtw_wave_cta<32><<<wgrid, kWideWarps * 32, 0, s>>>(L, order_dev, n_order, work_counter); break;
	default: return cudaErrorInvalidValue;
	}
	return cudaGetLastError();
}

} // namespace ra
