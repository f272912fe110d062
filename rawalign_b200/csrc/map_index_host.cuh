// map_index_host.cuh -- host side of the device index build and its export (included by map.cu only).
#pragma once

extern "C" int ra_index_build(ra_ctx *ctx, int32_t e, int32_t q, int32_t lq, int32_t k, uint32_t n_seq, const uint8_t *const *seq, const uint32_t *seq_len,
                              const float *pore_vals, ra_index **out)
{
	using namespace mapk;
	if (!ctx || !seq || !seq_len || !pore_vals || !out) return ra__fail(RA_EINVAL, "NULL argument");
	if (e < 2 || e > 9 || (lq + 2) * e > 63 || q < 1 || q > 32 || lq < 0 || lq > q || lq > 6 || k < 1 || k > 12) return ra__fail(RA_EINVAL, "bad index parameters e=%d q=%d lq=%d k=%d", e, q, lq, k);
	if (n_seq == 0 || n_seq > kMaxSeq) return ra__fail(RA_EUNSUP, "%u reference sequences (supported: 1..%u)", n_seq, kMaxSeq);
	uint64_t n_bases = 0, n_events = 0;
	uint32_t max_len = 0;
	for (uint32_t i = 0; i < n_seq; ++i) {
		if (seq_len[i] < (uint32_t)k) return ra__fail(RA_EINVAL, "sequence %u is shorter than the k-mer (%u < %d)", i, seq_len[i], k);
		n_bases += seq_len[i]; n_events += 2ull * (seq_len[i] - (uint32_t)k + 1); max_len = std::max(max_len, seq_len[i]);
	}
	if (n_events >= (1ull << 31)) return ra__fail(RA_EUNSUP, "reference too long for one index (%llu k-mers)", (unsigned long long)n_events);
	ra_mapopt_t o; int rc;
	if ((rc = ra_get_opt(ctx, &o))) return rc;
	MAP_CUDA(cudaSetDevice(ra__device(ctx)));
	ra_index *ix = new ra_index();
	ix->ctx = ctx; ix->w = 0; ix->e = e; ix->n = 0; ix->q = q; ix->lq = lq; ix->k = k; ix->n_seq = n_seq;
	ix->seq_len.assign(seq_len, seq_len + n_seq);
	auto bail = [&](int code) { ra_index_destroy(ix); return code; };
	uint64_t total = 0;
	std::vector<uint64_t> arena_off = arena_layout(ix, &total);
	if ((rc = ra_ref_alloc(ctx, total))) return bail(rc);
	float *arena = nullptr; uint64_t arena_n = 0;
	if ((rc = ra_ref_device(ctx, &arena, &arena_n))) return bail(rc);
	// jobs in the reference's order: per sequence the strand-flag-1 signal (forward_signals) then strand 0 (rawindex.cpp:141-147)
	const uint32_t n_jobs = 2 * n_seq;
	std::vector<StrandJob> jobs(n_jobs);
	std::vector<uint64_t> base_off(n_seq);
	uint64_t so = 0, bo = 0;
	for (uint32_t i = 0; i < n_seq; ++i) {
		base_off[i] = bo; bo += seq_len[i];
		for (int j = 0; j < 2; ++j) {
			StrandJob &J = jobs[2 * i + j];
			J.sig_off = arena_off[2 * i + j]; J.len = seq_len[i]; J.n_sig = seq_len[i] - (uint32_t)k + 1; J.seq = i; J.strand = j == 0 ? 1u : 0u; J.seed_off = so;
			so += J.n_sig;
		}
	}
	Buf d_bases, d_boff, d_jobs, d_pore, d_ms, d_hash, d_y, d_nseed, d_slot, d_cnt, d_off, d_cursor, d_nk;
	Buf *tmp[] = { &d_bases, &d_boff, &d_jobs, &d_pore, &d_ms, &d_hash, &d_y, &d_nseed, &d_slot, &d_cnt, &d_off, &d_cursor, &d_nk };
	auto cleanup = [&]() { for (Buf *b : tmp) b->release(); };
	auto fail_rc = [&](int code) { cleanup(); return bail(code); };
#define IB_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { cleanup(); ra_index_destroy(ix); return ra__fail(RA_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); } } while (0)
	const size_t n_pore = (size_t)1 << (2 * k);
	if ((rc = d_bases.reserve(n_bases)) || (rc = d_boff.reserve(n_seq * 8)) || (rc = d_jobs.reserve(n_jobs * sizeof(StrandJob))) || (rc = d_pore.reserve(n_pore * 4)) ||
	    (rc = d_ms.reserve(n_jobs * 16)) || (rc = d_hash.reserve(std::max<uint64_t>(so, 1) * 4)) || (rc = d_y.reserve(std::max<uint64_t>(so, 1) * 8)) ||
	    (rc = d_nseed.reserve(n_jobs * 4)) || (rc = d_nk.reserve(64))) return fail_rc(rc);
	for (uint32_t i = 0; i < n_seq; ++i) IB_CUDA(cudaMemcpy(d_bases.as<uint8_t>() + base_off[i], seq[i], seq_len[i], cudaMemcpyHostToDevice));
	IB_CUDA(cudaMemcpy(d_boff.p, base_off.data(), n_seq * 8, cudaMemcpyHostToDevice));
	IB_CUDA(cudaMemcpy(d_jobs.p, jobs.data(), n_jobs * sizeof(StrandJob), cudaMemcpyHostToDevice));
	IB_CUDA(cudaMemcpy(d_pore.p, pore_vals, n_pore * 4, cudaMemcpyHostToDevice));
	const dim3 wide((unsigned)std::min<uint64_t>(2048, ((uint64_t)max_len + 255) / 256), n_jobs);
	k_ib_levels<<<wide, 256>>>(d_bases.as<uint8_t>(), d_boff.as<uint64_t>(), d_jobs.as<StrandJob>(), n_jobs, d_pore.as<float>(), k, arena);
	if (o.fp_contract) k_ib_sums<true><<<n_jobs, 32>>>(d_jobs.as<StrandJob>(), n_jobs, arena, d_ms.as<double>());
	else k_ib_sums<false><<<n_jobs, 32>>>(d_jobs.as<StrandJob>(), n_jobs, arena, d_ms.as<double>());
	k_ib_normalise<<<wide, 256>>>(d_jobs.as<StrandJob>(), n_jobs, d_ms.as<double>(), arena);
	k_ib_sketch<<<n_jobs, 32>>>(d_jobs.as<StrandJob>(), n_jobs, arena, e, q, lq, d_hash.as<uint32_t>(), d_y.as<uint64_t>(), d_nseed.as<uint32_t>());
	std::vector<uint32_t> n_seeds(n_jobs);
	IB_CUDA(cudaMemcpy(n_seeds.data(), d_nseed.p, n_jobs * 4, cudaMemcpyDeviceToHost));
	uint64_t n_vals = 0;
	for (uint32_t v : n_seeds) n_vals += v;
	uint32_t bits = 4;
	while ((1ull << bits) < 2 * n_vals + 16) ++bits;               // >= 2 x the number of distinct keys, whatever it turns out to be
	if (bits > 31) return fail_rc(ra__fail(RA_EUNSUP, "index has too many seeds (%llu)", (unsigned long long)n_vals));
	const uint64_t n_slots = 1ull << bits;
	ix->slot_bits = bits; ix->n_vals = n_vals;
	if ((rc = ix->slots.reserve(n_slots * sizeof(Slot))) || (rc = ix->vals.reserve(std::max<uint64_t>(n_vals, 1) * 8)) || (rc = d_slot.reserve(std::max<uint64_t>(so, 1) * 4)) ||
	    (rc = d_cnt.reserve(n_slots * 4)) || (rc = d_off.reserve((n_slots + 1) * 8)) || (rc = d_cursor.reserve(n_slots * 4))) return fail_rc(rc);
	IB_CUDA(cudaMemset(ix->slots.p, 0, n_slots * sizeof(Slot)));
	IB_CUDA(cudaMemset(d_nk.p, 0, 8));
	const dim3 sgrid((unsigned)std::min<uint64_t>(2048, ((uint64_t)max_len + 255) / 256), n_jobs);
	k_ib_count<<<sgrid, 256>>>(d_jobs.as<StrandJob>(), d_nseed.as<uint32_t>(), n_jobs, d_hash.as<uint32_t>(), ix->slots.as<Slot>(), bits, d_slot.as<uint32_t>());
	k_ib_slot_counts<<<(unsigned)((n_slots + 255) / 256), 256>>>(ix->slots.as<Slot>(), (uint32_t)n_slots, d_cnt.as<uint32_t>());
	IB_CUDA(ra::launch_exclusive_scan_u32(d_cnt.as<uint32_t>(), d_off.as<uint64_t>(), (uint32_t)n_slots, 0));
	k_ib_slot_offsets<<<(unsigned)((n_slots + 255) / 256), 256>>>(ix->slots.as<Slot>(), (uint32_t)n_slots, d_off.as<uint64_t>(), d_cursor.as<uint32_t>(),
	                                                              d_nk.as<unsigned long long>());
	k_ib_scatter<<<sgrid, 256>>>(d_jobs.as<StrandJob>(), d_nseed.as<uint32_t>(), n_jobs, d_y.as<uint64_t>(), d_slot.as<uint32_t>(), ix->slots.as<Slot>(), d_cursor.as<uint32_t>(),
	                             ix->vals.as<uint64_t>());
	unsigned long long nk = 0;
	IB_CUDA(cudaMemcpy(&nk, d_nk.p, 8, cudaMemcpyDeviceToHost));   // also synchronises with the kernels above
	IB_CUDA(cudaGetLastError());
	ix->n_keys = nk;
	ra__add_launches(ctx, 9);
	cleanup();
#undef IB_CUDA
	if ((rc = finish_index_geometry(ctx, ix, arena_off))) return bail(rc);
	*out = ix;
	return RA_OK;
}

// The index as the flat table of ra_index_load_flat (keys ascending, positions of a key ascending = ri_idx_get's order), for
// comparison with the reference's index and for writing it back in the reference's file format.  Sizing call: keys == NULL.
extern "C" int ra_index_export(ra_index *ix, uint64_t *n_keys, uint64_t *n_vals, uint64_t *keys, uint64_t *key_off, uint64_t *vals)
{
	if (!ix || !n_keys || !n_vals) return ra__fail(RA_EINVAL, "NULL argument");
	*n_keys = ix->n_keys; *n_vals = ix->n_vals;
	if (!keys) return RA_OK;
	if (!key_off || !vals) return ra__fail(RA_EINVAL, "NULL argument");
	MAP_CUDA(cudaSetDevice(ra__device(ix->ctx)));
	const uint64_t n_slots = 1ull << ix->slot_bits;
	std::vector<Slot> slots(n_slots);
	std::vector<uint64_t> v(ix->n_vals);
	MAP_CUDA(cudaMemcpy(slots.data(), ix->slots.p, n_slots * sizeof(Slot), cudaMemcpyDeviceToHost));
	if (ix->n_vals) MAP_CUDA(cudaMemcpy(v.data(), ix->vals.p, ix->n_vals * 8, cudaMemcpyDeviceToHost));
	std::vector<uint64_t> used;                                    // key << 32 | slot: sorts by key without touching the table
	used.reserve(ix->n_keys);
	for (uint64_t h = 0; h < n_slots; ++h) if (slots[h].cnt) used.push_back(((uint64_t)slots[h].key << 32) | h);
	if (used.size() != ix->n_keys) return ra__fail(RA_EINVAL, "index table holds %zu keys, %llu expected", used.size(), (unsigned long long)ix->n_keys);
	std::sort(used.begin(), used.end());
	uint64_t w = 0;
	for (size_t i = 0; i < used.size(); ++i) {
		const Slot &s = slots[(uint32_t)used[i]];
		keys[i] = s.key; key_off[i] = w;
		std::copy(v.begin() + s.off, v.begin() + s.off + s.cnt, vals + w);
		std::sort(vals + w, vals + w + s.cnt);
		w += s.cnt;
	}
	key_off[used.size()] = w;
	return RA_OK;
}
