// dtw.cu -- K2/K3/K4 generic path: one warp per DTW problem, anti-diagonal wavefront.
//
// Semantics (bit-identical to the reference for every schedule, because each cell is
// min(min(top,left),topleft) + |a-b| -- min is exact, one rounding per cell):
//   banded : the cell set of DTW_global_slantedbanded_antidiagonalwise, src/dtw.cpp:273-520
//            (closed form in oracle/ra_oracle.c:ora_dtw_banded and DESIGN.md), out-of-set = 1e10f;
//   full   : DTW_global, src/dtw.cpp:37-66 (whole rectangle);
//   tb     : DTW_global_tb, src/dtw.cpp:595-667 -- full rectangle, 2-bit direction per cell packed
//            16 cells per 32-bit word in HBM, then the tie rule of :629-653 walked on the device.
//
// This is the shape-agnostic kernel: lanes stride over the rows of one anti-diagonal, three rolling
// anti-diagonals live in a per-warp shared-memory ring (or in an HBM/L2 ring when a band is wider than
// the ring), warps pull problems from a global work counter.  k_dtw_tiny (below) takes the sparse anchor-to-anchor shapes
// (R' <= 2) one per thread and dtw_wave.cu the problems whose band fits a register wavefront; the rest lands here.
#include "dtw_core.cuh"

namespace ra {

constexpr int kWarpsPerCta = 4;
constexpr uint32_t kSmemRingCap = 1024;          // floats per anti-diagonal ring in shared memory

uint32_t dtw_generic_smem_ring_cap() { return kSmemRingCap; }

template <bool TB>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_dtw_generic(DtwLaunch L)
{
	extern __shared__ float smem_ring[];           // [kWarpsPerCta][3][kSmemRingCap]
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const uint32_t gwarp = blockIdx.x * kWarpsPerCta + wid;
	float *const my_smem = smem_ring + (size_t)wid * 3 * kSmemRingCap;
	float *const my_gmem = L.ring_scratch ? L.ring_scratch + (size_t)gwarp * 3 * L.ring_cap : nullptr;

	for (;;) {
		uint32_t slot = 0;
		if (lane == 0) slot = atomicAdd(L.work_counter, 1u);
		slot = __shfl_sync(0xffffffffu, slot, 0);
		if (slot >= L.n_items) break;
		const uint32_t k = L.order ? L.order[slot] : slot;
		const ra_dtw_item_t it = L.items[k];
		const float *A = L.a_arena + it.a_off, *B = L.b_arena + it.b_off;
		int n = (int)it.a_len, m = (int)it.b_len;
		int band = TB ? -1 : it.band_radius;
		if (!TB && n < m) { const float *t = A; A = B; B = t; int tl = n; n = m; m = tl; }   // dtw.cpp:284-292 (harmless for full)
		if (n == 0 || m == 0) { if (lane == 0) { L.cost[k] = kInf; if (TB) L.path_len[k] = 0; } continue; }

		DiagIter di; di.init(n, m, band);
		const int width = di.max_width();
		uint32_t cap = 1; while ((int)cap < width) cap <<= 1;
		float *ring = (cap <= kSmemRingCap) ? my_smem : my_gmem;
		if (cap <= kSmemRingCap) cap = kSmemRingCap;
		else cap = L.ring_cap;
		const uint32_t mask = cap - 1;
		float *r0 = ring, *r1 = ring + cap, *r2 = ring + 2 * (size_t)cap;
		int lo1 = 0, hi1 = -1, lo2 = 0, hi2 = -1;

		uint32_t *dirs = nullptr; uint32_t dstride = 0;
		if (TB) { dirs = L.dirs + L.dir_off[k]; dstride = ((uint32_t)m + 15u) >> 4; }

		// diagonal 0: the corner
		if (lane == 0) r0[0] = fabsf(__ldg(A) - __ldg(B));
		__syncwarp();
		int lo0 = 0, hi0 = 0;

		while (true) {
			// rotate r2 <- r1 <- r0 <- old r2
			float *t = r2; r2 = r1; r1 = r0; r0 = t;
			lo2 = lo1; hi2 = hi1; lo1 = lo0; hi1 = hi0;
			if (!di.next()) break;
			lo0 = di.lo; hi0 = di.hi;
			const int d = di.d;
			for (int jb = lo0 & ~31; jb <= hi0; jb += 32) {
				const int j = jb + lane;
				const bool act = (j >= lo0) && (j <= hi0);
				uint32_t code = 0;
				if (act) {
					const int i = d - j;
					// (i, j-1) and (i-1, j) on diagonal d-1; (i-1, j-1) on diagonal d-2
					const float top = (j - 1 >= lo1 && j - 1 <= hi1) ? r1[(uint32_t)(j - 1) & mask] : kInf;
					const float left = (j >= lo1 && j <= hi1) ? r1[(uint32_t)j & mask] : kInf;
					const float tl = (j - 1 >= lo2 && j - 1 <= hi2) ? r2[(uint32_t)(j - 1) & mask] : kInf;
					const float v = fminf(fminf(top, left), tl) + fabsf(__ldg(A + i) - __ldg(B + j));
					r0[(uint32_t)j & mask] = v;
					if (TB) {
						// dtw.cpp:636-653 with left = dp[i-1][j], top = dp[i][j-1], topleft = dp[i-1][j-1]
						if (left < fminf(top, tl)) code = 1u;           // i--
						else if (top < fminf(left, tl)) code = 2u;      // j--
						else code = 0u;                                 // i--, j--
					}
				}
				if (TB) {
					const uint32_t bits = act ? (code << (2 * (j & 15))) : 0u;
					const uint32_t w_lo = __reduce_or_sync(0xffffffffu, lane < 16 ? bits : 0u);
					const uint32_t w_hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? bits : 0u);
					const uint32_t wbase = (uint32_t)jb >> 4;
					if (lane == 0 && wbase < dstride) dirs[(size_t)d * dstride + wbase] = w_lo;
					if (lane == 16 && wbase + 1 < dstride) dirs[(size_t)d * dstride + wbase + 1] = w_hi;
				}
			}
			__syncwarp();
		}
		// after the final rotation the last diagonal is r1
		if (lane == 0) {
			float res = r1[(uint32_t)(m - 1) & mask];
			if (it.exclude_last) res -= fabsf(__ldg(A + n - 1) - __ldg(B + m - 1));   // dtw.cpp:514-516 / :60-62
			L.cost[k] = res;
		}
		if (TB) {
			__syncwarp();
			if (lane == 0) {
				// dtw.cpp:616-654; written back-to-front so the slot ends in forward order
				const uint64_t so = L.slot_off[k];
				const uint32_t capk = (uint32_t)n + (uint32_t)m - 1u;
				uint32_t i = (uint32_t)n - 1u, j = (uint32_t)m - 1u, w = capk, len = 0;
				for (;;) {
					--w; ++len;
					L.path_i[so + w] = i; L.path_j[so + w] = j;
					if (L.path_diff) L.path_diff[so + w] = fabsf(__ldg(A + i) - __ldg(B + j));
					if (i == 0 && j == 0) break;
					if (i == 0) --j;
					else if (j == 0) --i;
					else {
						const uint32_t word = __ldcg(dirs + (size_t)(i + j) * dstride + (j >> 4));   // L2 read: other lanes wrote it
						const uint32_t code = (word >> (2 * (j & 15))) & 3u;
						if (code == 1u) --i;
						else if (code == 2u) --j;
						else { --i; --j; }
					}
				}
				L.path_len[k] = it.exclude_last ? len - 1u : len;         // dtw.cpp:659-663 drops the final element
			}
		}
		__syncwarp();
	}
}

// Copies each item's path from its (end-aligned) slot to the CSR output.
__global__ void __launch_bounds__(256)
k_path_compact(const uint32_t *__restrict__ slot_i, const uint32_t *__restrict__ slot_j, const float *__restrict__ slot_d,
               const uint64_t *__restrict__ slot_off, const ra_dtw_item_t *__restrict__ items,
               const uint32_t *__restrict__ path_len, const uint64_t *__restrict__ path_off, uint32_t n_items,
               uint32_t *__restrict__ out_i, uint32_t *__restrict__ out_j, float *__restrict__ out_d)
{
	const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (k >= n_items) return;
	const uint32_t len = path_len[k];
	const uint32_t full_len = len + (items[k].exclude_last ? 1u : 0u);   // the slot still holds the dropped final element
	const uint32_t capk = items[k].a_len + items[k].b_len - 1u;
	const uint64_t src = slot_off[k] + (capk - full_len);
	const uint64_t dst = path_off[k];
	for (uint32_t t = lane; t < len; t += 32) {
		out_i[dst + t] = slot_i[src + t];
		out_j[dst + t] = slot_j[src + t];
		if (out_d) out_d[dst + t] = slot_d[src + t];
	}
}

// Items whose slant-extended radius R' is 1 or 2 (band_radius 1: the sparse anchor-to-anchor shapes): the band is at most
// five cells wide, so a wavefront across lanes would leave most of a warp idle.  One item per THREAD through the
// register-resident anti-diagonal evaluator of the chain kernels (R' = 1 embedded in the R' = 2 schedule, dtw_core.cuh);
// the host sorts these items by height, so the lanes of a warp run equally long.
__global__ void __launch_bounds__(128)
k_dtw_tiny(DtwLaunch L, const uint32_t *__restrict__ order, uint32_t n_order)
{
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n_order) return;
	const uint32_t k = order[t];
	const ra_dtw_item_t it = L.items[k];
	const float *a = L.a_arena + it.a_off, *b = L.b_arena + it.b_off;
	const bool swap = it.a_len < it.b_len;                               // dtw.cpp:284-292: A is the longer sequence
	const LdgLoader LA{swap ? b : a}, LB{swap ? a : b};
	const int n = (int)(swap ? it.b_len : it.a_len), m = (int)(swap ? it.a_len : it.b_len);
	const int rp = slant_radius(n, m, it.band_radius);
	float c = thread_dtw_diag<2, false, true>(LA, n, LB, m, nullptr, rp == 1);
	if (it.exclude_last) c -= fabsf(__ldg(a + it.a_len - 1) - __ldg(b + it.b_len - 1));   // dtw.cpp:514-516
	L.cost[k] = c;
}

bool dtw_tiny_supported(uint32_t a_len, uint32_t b_len, int32_t band_radius)
{
	if (band_radius < 1 || band_radius > 2 || a_len == 0 || b_len == 0) return false;
	const uint64_t n = a_len > b_len ? a_len : b_len, m = a_len > b_len ? b_len : a_len;
	if (n >= (1u << 28)) return false;
	const uint64_t rp = (uint64_t)band_radius + ((n - m) * (uint64_t)band_radius + n - 1) / n;    // dtw.cpp:298-300
	return rp <= 2;
}

cudaError_t launch_dtw_tiny(const DtwLaunch &L, const uint32_t *order_dev, uint32_t n_order, cudaStream_t s)
{
	if (n_order == 0) return cudaSuccess;
	k_dtw_tiny<<<(n_order + 127) / 128, 128, 0, s>>>(L, order_dev, n_order);
	return cudaGetLastError();
}

cudaError_t launch_dtw_generic(const DtwLaunch &L, bool traceback, int n_sm, cudaStream_t s, int *n_launches)
{
	if (L.n_items == 0) return cudaSuccess;
	cudaError_t e = cudaMemsetAsync(L.work_counter, 0, sizeof(uint32_t), s);
	if (e != cudaSuccess) return e;
	const size_t smem = (size_t)kWarpsPerCta * 3 * kSmemRingCap * sizeof(float);
	static unsigned long long attr_devices = 0;                       // the opt-in is per function AND per device
	int dev = 0;
	cudaGetDevice(&dev);
	if (!((attr_devices >> (dev & 63)) & 1ull)) {
		e = cudaFuncSetAttribute(k_dtw_generic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e == cudaSuccess) e = cudaFuncSetAttribute(k_dtw_generic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) return e;
		__atomic_fetch_or(&attr_devices, 1ull << (dev & 63), __ATOMIC_RELAXED);
	}
	// persistent grid: 4 CTAs of 4 warps per SM (shared-memory bound), never more warps than problems
	uint32_t ctas = (uint32_t)n_sm * 4u;
	const uint32_t need = (L.n_items + kWarpsPerCta - 1) / kWarpsPerCta;
	if (ctas > need) ctas = need;
	if (traceback) k_dtw_generic<true><<<ctas, kWarpsPerCta * 32, smem, s>>>(L);
	else k_dtw_generic<false><<<ctas, kWarpsPerCta * 32, smem, s>>>(L);
	if (n_launches) *n_launches += 1;
	return cudaGetLastError();
}

cudaError_t launch_path_compact(const uint32_t *slot_i, const uint32_t *slot_j, const float *slot_d,
                                const uint64_t *slot_off, const ra_dtw_item_t *items, const uint32_t *path_len,
                                const uint64_t *path_off, uint32_t n_items, uint32_t *out_i, uint32_t *out_j,
                                float *out_d, cudaStream_t s)
{
	if (n_items == 0) return cudaSuccess;
	const int threads = 256;
	const unsigned blocks = (unsigned)(((uint64_t)n_items * 32 + threads - 1) / threads);
	k_path_compact<<<blocks, threads, 0, s>>>(slot_i, slot_j, slot_d, slot_off, items, path_len, path_off, n_items, out_i, out_j, out_d);
	return cudaGetLastError();
}

} // namespace ra
