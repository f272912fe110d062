/*
 * ref_shim.cpp -- thin C entry points around the UNMODIFIED reference objects.  TEST INFRASTRUCTURE.
 *
 * oracle/Makefile compiles /root/reference/src/{dtw.cpp,revent.c,roptions.c,kalloc.c} where they
 * lie (nothing is copied into this repository) with the reference's own flags and links them with
 * this file into oracle/_ref/libra_ref*.so.  The shim adds no arithmetic: it only unpacks
 * arguments, calls the reference function, and copies the result out, so Python (ctypes) and
 * bench.py --impl reference can drive the real reference implementation of the hot path.
 */
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "dtw.hpp"       /* -I/root/reference/src */
#include "revent.h"
#include "kalloc.h"

extern "C" {

void ref_mapopt_init(ri_mapopt_t *opt) { ri_mapopt_init(opt); }
uint32_t ref_sizeof_mapopt(void) { return (uint32_t)sizeof(ri_mapopt_t); }

/* detect_events with the five detector scalars overriding the defaults. km = NULL -> malloc. */
uint32_t ref_detect_events(const float *sig, uint32_t s_len, uint32_t w1, uint32_t w2, float thr1, float thr2,
                           float peak_height, float *ev_out)
{
	ri_mapopt_t opt; ri_mapopt_init(&opt);
	opt.window_length1 = w1; opt.window_length2 = w2; opt.threshold1 = thr1; opt.threshold2 = thr2;
	opt.peak_height = peak_height;
	uint32_t n = 0;
	float *ev = detect_events(0, s_len, sig, &opt, &n);
	if (ev) { memcpy(ev_out, ev, (size_t)n * sizeof(float)); ri_kfree(0, ev); }
	return ev ? n : 0;
}

void ref_detect_events_batch(const float *sig, const uint64_t *chunk_off, const uint32_t *chunk_len,
                             uint32_t n_chunks, float *ev_out, uint64_t *ev_off)
{
	ri_mapopt_t opt; ri_mapopt_init(&opt);
	void *km = ri_km_init();
	uint64_t o = 0;
	for (uint32_t c = 0; c < n_chunks; ++c) {
		ev_off[c] = o;
		uint32_t n = 0;
		float *ev = detect_events(km, chunk_len[c], sig + chunk_off[c], &opt, &n);
		if (ev) { memcpy(ev_out + o, ev, (size_t)n * sizeof(float)); ri_kfree(km, ev); o += n; }
	}
	ev_off[n_chunks] = o;
	ri_km_destroy(km);
}

float ref_dtw_global(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int excl)
{ return DTW_global(a, a_len, b, b_len, excl != 0); }

float ref_dtw_global_slow(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int excl)
{ return DTW_global_slow(a, a_len, b, b_len, excl != 0); }

float ref_dtw_banded(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int R, int excl)
{ return DTW_global_slantedbanded_antidiagonalwise(a, a_len, b, b_len, R, excl != 0); }

float ref_dtw_slantedbanded(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int R, int excl)
{ return DTW_global_slantedbanded(a, a_len, b, b_len, R, excl != 0); }

uint32_t ref_dtw_global_tb(const float *a, uint32_t a_len, const float *b, uint32_t b_len, int excl,
                           float *cost, uint32_t *pi, uint32_t *pj, float *pd)
{
	dtw_result r = DTW_global_tb(a, a_len, b, b_len, excl != 0);
	*cost = r.cost;
	for (size_t k = 0; k < r.alignment.size(); ++k) {
		pi[k] = (uint32_t)r.alignment[k].position.i;
		pj[k] = (uint32_t)r.alignment[k].position.j;
		if (pd) pd[k] = r.alignment[k].difference;
	}
	return (uint32_t)r.alignment.size();
}

typedef struct { uint64_t a_off, b_off; uint32_t a_len, b_len; int32_t band_radius; uint32_t exclude_last; } shim_item_t;

void ref_dtw_batch(const float *a_arena, const float *b_arena, const shim_item_t *items, uint32_t n_items, float *cost)
{
	for (uint32_t k = 0; k < n_items; ++k) {
		const shim_item_t &it = items[k];
		cost[k] = it.band_radius < 0
			? DTW_global(a_arena + it.a_off, it.a_len, b_arena + it.b_off, it.b_len, it.exclude_last != 0)
			: DTW_global_slantedbanded_antidiagonalwise(a_arena + it.a_off, it.a_len, b_arena + it.b_off, it.b_len,
			                                            it.band_radius, it.exclude_last != 0);
	}
}

/* The input generator of the reference's own unit test (check_dtw.cpp:13-27 draws
 * uniform_real_distribution<double>(-2.5,2.5) from mt19937(seed) and :43-52 narrows to float);
 * written here with <random> so fixtures use exactly the shapes/seeds check_dtw uses. */
void ref_checkdtw_vector(int n, unsigned seed, float *out)
{
	std::mt19937 rng(seed);
	std::uniform_real_distribution<double> dist(-2.5, 2.5);
	for (int i = 0; i < n; ++i) out[i] = (float)dist(rng);
}

} /* extern "C" */
