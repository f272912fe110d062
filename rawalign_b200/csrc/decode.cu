// decode.cu -- raw-signal decode on the device: the conversion loop of ri_read_sig,
// src/rsig.cpp:213-222, for a batch of reads.
//
//   float scale = ran / dig;                                   (host, passed per read)
//   pA = ((float)raw[i] + offset) * scale;                     two fp32 roundings, no contraction possible
//   kept iff 30 < pA < 200, compacted in order                 -> sig[0 .. l_sig)
//
// Shipping the DAC samples as int16 halves the host->device bytes of the path (2 B instead of 4 B per sample)
// and removes a host pass.  One warp per read: 256 samples per iteration (8 per lane, one 16-byte load when the
// read is 16-byte aligned), per-lane keep masks, warp scan of the keep counts, compaction through shared memory
// so that the float arena is written with coalesced 128-byte rows.  The output of read r starts at the same
// element offset as its input (sig[raw_off[r]]), so a read's chunk c is sig[raw_off[r] + c*chunk_size ...].
// HBM-bound: 2 B in + <= 4 B out per sample.
#include "ra_common.cuh"

namespace ra {

constexpr int kDecWarps = 4;

__global__ void __launch_bounds__(kDecWarps * 32)
k_decode_raw(const int16_t *__restrict__ raw, const uint64_t *__restrict__ raw_off, const uint32_t *__restrict__ raw_len,
             const float *__restrict__ offset, const float *__restrict__ scale, uint32_t n_reads, float *__restrict__ sig,
             uint32_t *__restrict__ l_sig)
{
	__shared__ float stage[kDecWarps][256];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const uint32_t r = blockIdx.x * kDecWarps + (uint32_t)wid;
	if (r >= n_reads) return;
	const uint64_t off = raw_off[r];
	const uint32_t len = raw_len[r];
	const float o = offset[r], sc = scale[r];
	const int16_t *__restrict__ in = raw + off;
	float *__restrict__ out = sig + off;
	float *st = stage[wid];
	const bool vec = ((reinterpret_cast<uintptr_t>(in) & 15u) == 0);
	uint32_t kept = 0;
	// 16-byte load of the lane's 8 samples of the tile at `base` (zeros past the end); issued one tile ahead
	auto fetch = [&](uint32_t base) -> uint4 {
		const uint32_t p0 = base + 8u * (uint32_t)lane;
		if (vec && p0 + 8u <= len) return __ldcs(reinterpret_cast<const uint4 *>(in + p0));
		uint32_t w[4] = { 0u, 0u, 0u, 0u };
		for (int e = 0; e < 8; ++e)
			if (p0 + (uint32_t)e < len) w[e >> 1] |= (uint32_t)(uint16_t)in[p0 + e] << (16 * (e & 1));
		return make_uint4(w[0], w[1], w[2], w[3]);
	};
	uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
	if (len) nxt = fetch(0u);
	for (uint32_t base = 0; base < len; base += 256u) {
		const uint32_t p0 = base + 8u * (uint32_t)lane;
		const uint4 q = nxt;
		if (base + 256u < len) nxt = fetch(base + 256u);
		int16_t v[8];
		{
			const uint32_t w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
			for (int e = 0; e < 4; ++e) { v[2 * e] = (int16_t)(w[e] & 0xffffu); v[2 * e + 1] = (int16_t)(w[e] >> 16); }
		}
		float pa[8];
		uint32_t mask = 0;
#pragma unroll
		for (int e = 0; e < 8; ++e) {
			pa[e] = __fmul_rn(__fadd_rn((float)v[e], o), sc);                       // rsig.cpp:219
			if (p0 + (uint32_t)e < len && pa[e] > 30.0f && pa[e] < 200.0f) mask |= 1u << e;   // :220
		}
		const uint32_t cnt = (uint32_t)__popc(mask);
		uint32_t incl = cnt;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += y; }
		const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
		uint32_t w = incl - cnt;
#pragma unroll
		for (int e = 0; e < 8; ++e)
			if (mask & (1u << e)) st[w++] = pa[e];
		__syncwarp();
		for (uint32_t t = lane; t < total; t += 32u) out[kept + t] = st[t];
		__syncwarp();
		kept += total;
	}
	if (lane == 0) l_sig[r] = kept;
}

// Chunk tables travel host -> device by a kernel reading the pinned (device-mapped) staging buffer instead of a copy-engine
// transfer: the H2D engine's queue is full of signal slices, and a table queued behind them would hold back K1.
__global__ void __launch_bounds__(256)
k_fetch_tables(const uint64_t *__restrict__ off_host, const uint32_t *__restrict__ len_host, uint64_t *__restrict__ off_dev,
               uint32_t *__restrict__ len_dev, uint32_t n)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) { off_dev[i] = off_host[i]; len_dev[i] = len_host[i]; }
}

cudaError_t launch_fetch_tables(const uint64_t *off_host, const uint32_t *len_host, uint64_t *off_dev, uint32_t *len_dev, uint32_t n, cudaStream_t s)
{
	if (n == 0) return cudaSuccess;
	k_fetch_tables<<<(n + 255) / 256, 256, 0, s>>>(off_host, len_host, off_dev, len_dev, n);
	return cudaGetLastError();
}

cudaError_t launch_decode_raw(const int16_t *raw, const uint64_t *raw_off, const uint32_t *raw_len, const float *offset,
                              const float *scale, uint32_t n_reads, float *sig, uint32_t *l_sig, cudaStream_t s)
{
	if (n_reads == 0) return cudaSuccess;
	k_decode_raw<<<(n_reads + kDecWarps - 1) / kDecWarps, kDecWarps * 32, 0, s>>>(raw, raw_off, raw_len, offset, scale, n_reads, sig, l_sig);
	return cudaGetLastError();
}

} // namespace ra
