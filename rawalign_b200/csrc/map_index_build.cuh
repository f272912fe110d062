// map_index_build.cuh -- index construction on the device (included by map.cu only):
// ri_idx_gen's per-sequence work (src/rawindex.cpp:127-153: ri_seq_to_sig for strand flag 1 then 0, src/rsig.cpp:7-41, and
// ri_sketch of both signals, src/rsketch.c:223-274) and the grouping of worker_post (:194-247) as a device hash table.
//
//   k_ib_levels     one thread per k-mer position: the k-mer code of the strand's walk (:13-24) -> pore_vals[kmer]
//   k_ib_sums       one warp per (sequence, strand): sum and sum of squares in fp64 IN SEQUENCE ORDER (:29-31; the fused
//                   multiply-add of the reference build when fp_contract is set) -- the only order that is bit-exact, so
//                   the warp loads 32 values at a time and every lane replays the 32 additions on shuffled values
//   k_ib_normalise  (float)(((double)v - mean) / sd), parallel (:37-38); the tail of the len-float array stays 0
//   k_ib_sketch     one warp per (sequence, strand): the accept filter / quantisation / hash of ri_sketch_reg (the code of
//                   k_seed over a whole strand), seeds appended to the strand's slice of (hash, y) arrays
//   k_ib_count / k_ib_scatter   open-addressing table: claim-or-increment per seed, offsets from one exclusive scan over the
//                   slots, positions scattered behind a per-slot cursor.  The positions of one key end up contiguous but in
//                   no particular order: every consumer re-sorts the anchors it makes from them (rmap.cpp:396-401), and
//                   ra_index_export sorts them on the host into ri_idx_get's order for comparison with the reference.
#pragma once

namespace mapk {

struct StrandJob { uint64_t sig_off; uint32_t len, n_sig, seq, strand; uint64_t seed_off; };   // sig_off into the mirror arena; n_sig = len - k + 1 (0 if len < k)

__global__ void __launch_bounds__(256)
k_ib_levels(const uint8_t *__restrict__ bases, const uint64_t *__restrict__ base_off, const StrandJob *__restrict__ jobs, uint32_t n_jobs,
            const float *__restrict__ pore, int k, float *__restrict__ arena)
{
	const uint32_t jb = blockIdx.y;
	if (jb >= n_jobs) return;
	const StrandJob J = jobs[jb];
	const uint8_t *__restrict__ s = bases + base_off[J.seq];
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < J.len; j += (uint64_t)gridDim.x * blockDim.x) {
		float v = 0.0f;                                            // positions past len - k stay zero (rawindex.cpp:136-137 callocs len floats)
		if (j < J.n_sig) {
			uint32_t kmer = 0;
			for (int t = 0; t < k; ++t) {
				const uint32_t pos = J.strand ? (uint32_t)(J.len - 1 - (j + t)) : (uint32_t)(j + t);   // rsig.cpp:15-16
				const uint32_t c = s[pos];
				const uint32_t code = c < 4 ? (J.strand ? (3u ^ c) : c) : 0u;                          // :18-24
				kmer = (kmer << 2) | code;
			}
			v = pore[kmer];
		}
		arena[J.sig_off + j] = v;
	}
}

template <bool CONTRACT>
__global__ void __launch_bounds__(32)
k_ib_sums(const StrandJob *__restrict__ jobs, uint32_t n_jobs, const float *__restrict__ arena, double *__restrict__ mean_sd)
{
	const uint32_t jb = blockIdx.x;
	const int lane = threadIdx.x;
	if (jb >= n_jobs) return;
	const StrandJob J = jobs[jb];
	const float *__restrict__ x = arena + J.sig_off;
	double sum = 0.0, sum2 = 0.0;
	// the dependent chain is one DADD and one DFMA per value; everything else (load of the next block, shuffles, conversions) is
	// hoisted off it: full blocks run fully unrolled on 32 converted values held in registers
	float v = lane < J.n_sig ? x[lane] : 0.0f;
	for (uint32_t b = 0; b < J.n_sig; b += 32) {
		const uint32_t nb = b + 32;
		const float vn = nb + lane < J.n_sig ? x[nb + lane] : 0.0f;
		const uint32_t cnt = min(32u, J.n_sig - b);
		if (cnt == 32u) {
			double c[32];
#pragma unroll
			for (int t = 0; t < 32; ++t) c[t] = (double)__shfl_sync(0xffffffffu, v, t);
#pragma unroll
			for (int t = 0; t < 32; ++t) {
				sum = __dadd_rn(sum, c[t]);                                                          // rsig.cpp:30
				sum2 = CONTRACT ? __fma_rn(c[t], c[t], sum2) : __dadd_rn(sum2, __dmul_rn(c[t], c[t]));   // :31
			}
		} else {
			for (uint32_t t = 0; t < cnt; ++t) {
				const double c = (double)__shfl_sync(0xffffffffu, v, (int)t);
				sum = __dadd_rn(sum, c);
				sum2 = CONTRACT ? __fma_rn(c, c, sum2) : __dadd_rn(sum2, __dmul_rn(c, c));
			}
		}
		v = vn;
	}
	if (lane == 0) {
		const double n = (double)J.n_sig;
		const double mean = __ddiv_rn(sum, n), q = __ddiv_rn(sum2, n);                   // :34-35
		const double var = CONTRACT ? __fma_rn(-mean, mean, q) : __dsub_rn(q, __dmul_rn(mean, mean));
		mean_sd[2 * jb] = mean; mean_sd[2 * jb + 1] = __dsqrt_rn(var);
	}
}

__global__ void __launch_bounds__(256)
k_ib_normalise(const StrandJob *__restrict__ jobs, uint32_t n_jobs, const double *__restrict__ mean_sd, float *__restrict__ arena)
{
	const uint32_t jb = blockIdx.y;
	if (jb >= n_jobs) return;
	const StrandJob J = jobs[jb];
	const double mean = mean_sd[2 * jb], sd = mean_sd[2 * jb + 1];
	for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < J.n_sig; j += (uint64_t)gridDim.x * blockDim.x)
		arena[J.sig_off + j] = __double2float_rn(__ddiv_rn(__dsub_rn((double)arena[J.sig_off + j], mean), sd));   // :37-38
}

// ri_sketch_reg over a whole strand (the same warp scheme as k_seed; the symbols of the last e accepted events live in a
// small ring, because a strand is too long for shared memory)
__global__ void __launch_bounds__(32)
k_ib_sketch(const StrandJob *__restrict__ jobs, uint32_t n_jobs, const float *__restrict__ arena, int e, int q, int lq,
            uint32_t *__restrict__ seed_hash, uint64_t *__restrict__ seed_y, uint32_t *__restrict__ n_seeds)
{
	__shared__ uint8_t ring[64];                                   // symbols of the accepted events, index = rank & 63
	const uint32_t jb = blockIdx.x;
	const int lane = threadIdx.x;
	if (jb >= n_jobs) return;
	const StrandJob J = jobs[jb];
	const float *__restrict__ ev = arena + J.sig_off;
	const uint32_t quant_bit = (uint32_t)lq + 2u, shift_r = 32u - (uint32_t)q, mask_l_quant = (1u << lq) - 1u;
	const uint64_t mask = (1ull << 32) - 1, mask_events = (1ull << (quant_bit * (uint32_t)e)) - 1;
	const uint64_t id_shift = (uint64_t)J.seq << 32;
	float last = 0.0f;
	uint32_t n_acc = 0;
	float v = (uint32_t)lane < J.n_sig ? ev[lane] : 0.0f;
	for (uint32_t b = 0; b < J.n_sig; b += 32) {
		const uint32_t idx = b + lane;
		const float vn = idx + 32 < J.n_sig ? ev[idx + 32] : 0.0f;  // next block's load is in flight during this block's chain
		const uint32_t cnt = min(32u, J.n_sig - b);
		uint32_t accmask = 0;
		if (cnt == 32u && b > 0) {                                  // full block: shuffles hoisted, the chain is subtract / compare / select
			float xs[32];
#pragma unroll
			for (int t = 0; t < 32; ++t) xs[t] = __shfl_sync(0xffffffffu, v, t);
#pragma unroll
			for (int t = 0; t < 32; ++t) {
				const bool rej = fabsf(__fsub_rn(xs[t], last)) < 0.3f || xs[t] == 3.402823466e+32f;            // rsketch.c:243
				last = rej ? last : xs[t];
				accmask |= rej ? 0u : (1u << t);
			}
		} else {
			for (uint32_t t = 0; t < cnt; ++t) {
				const float x = __shfl_sync(0xffffffffu, v, (int)t);
				const bool rej = ((b + t) > 0 && fabsf(__fsub_rn(x, last)) < 0.3f) || x == 3.402823466e+32f;   // rsketch.c:243
				if (!rej) { last = x; accmask |= 1u << t; }
			}
		}
		const bool mine = (accmask >> lane) & 1u;
		const uint32_t rank = n_acc + __popc(accmask & ((1u << lane) - 1u));
		// the e symbols a lane needs are at most e - 1 ranks behind its own: earlier ranks of this step come from the lanes
		// below it, older ones from the ring (written by previous steps)
		const uint32_t sbits = __float_as_uint(v);
		const uint32_t sym = mine ? ((sbits >> 30 << lq) | ((sbits >> shift_r) & mask_l_quant)) : 0u;   // :247
		__syncwarp();
		if (mine) ring[rank & 63u] = (uint8_t)sym;
		__syncwarp();
		if (mine && rank + 1u >= (uint32_t)e) {
			uint64_t qv = 0;
			for (int j = e - 1; j >= 0; --j) qv = (qv << quant_bit) | ring[(rank - (uint32_t)j) & 63u];   // :254
			qv &= mask_events;
			const uint64_t o = J.seed_off + (rank + 1u - (uint32_t)e);
			seed_hash[o] = (uint32_t)hash64(qv, mask);                                                 // :255 (x >> RI_HASH_SHIFT)
			seed_y[o] = id_shift | ((uint64_t)idx << 1) | (uint64_t)J.strand;                          // :253
		}
		n_acc += __popc(accmask);
		__syncwarp();
		v = vn;
	}
	if (lane == 0) n_seeds[jb] = n_acc + 1u >= (uint32_t)e ? n_acc + 1u - (uint32_t)e : 0u;
}

// slot word = key | count << 32 (count == 0: empty)
__global__ void __launch_bounds__(256)
k_ib_count(const StrandJob *__restrict__ jobs, const uint32_t *__restrict__ n_seeds, uint32_t n_jobs, const uint32_t *__restrict__ seed_hash,
           Slot *__restrict__ slots, uint32_t slot_bits, uint32_t *__restrict__ seed_slot)
{
	const uint32_t jb = blockIdx.y;
	if (jb >= n_jobs) return;
	const uint64_t base = jobs[jb].seed_off;
	const uint32_t n = n_seeds[jb], m = (1u << slot_bits) - 1u;
	unsigned long long *w = reinterpret_cast<unsigned long long *>(slots);   // 16-byte slots: word 2 * h
	for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t key = seed_hash[base + s];
		uint32_t h = (uint32_t)(key * 0x9E3779B1u) >> (32 - slot_bits);
		for (;;) {
			unsigned long long cur = w[2ull * h];
			if ((cur >> 32) == 0ull) {
				const unsigned long long prev = atomicCAS(&w[2ull * h], 0ull, (unsigned long long)key | (1ull << 32));
				if (prev == 0ull) break;
				cur = prev;
			}
			if ((uint32_t)cur == key) { atomicAdd(&w[2ull * h], 1ull << 32); break; }
			h = (h + 1u) & m;
		}
		seed_slot[base + s] = h;
	}
}

__global__ void __launch_bounds__(256)
k_ib_slot_counts(const Slot *__restrict__ slots, uint32_t n_slots, uint32_t *__restrict__ cnt)
{
	const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
	if (h < n_slots) cnt[h] = slots[h].cnt;
}

__global__ void __launch_bounds__(256)
k_ib_slot_offsets(Slot *__restrict__ slots, uint32_t n_slots, const uint64_t *__restrict__ off, uint32_t *__restrict__ cursor, unsigned long long *__restrict__ n_keys)
{
	const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
	if (h >= n_slots) return;
	slots[h].off = off[h]; cursor[h] = 0;
	const unsigned used = __ballot_sync(__activemask(), slots[h].cnt != 0);
	if ((threadIdx.x & 31) == 0 && used) atomicAdd(n_keys, (unsigned long long)__popc(used));
}

__global__ void __launch_bounds__(256)
k_ib_scatter(const StrandJob *__restrict__ jobs, const uint32_t *__restrict__ n_seeds, uint32_t n_jobs, const uint64_t *__restrict__ seed_y,
             const uint32_t *__restrict__ seed_slot, const Slot *__restrict__ slots, uint32_t *__restrict__ cursor, uint64_t *__restrict__ vals)
{
	const uint32_t jb = blockIdx.y;
	if (jb >= n_jobs) return;
	const uint64_t base = jobs[jb].seed_off;
	const uint32_t n = n_seeds[jb];
	for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t h = seed_slot[base + s];
		vals[slots[h].off + atomicAdd(&cursor[h], 1u)] = seed_y[base + s];
	}
}

} // namespace mapk
