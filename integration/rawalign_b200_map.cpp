// rawalign_b200_map.cpp -- a compiled C++ consumer of the C ABI (include/rawalign_b200.h): the map phase of `rawalign`
// for a batch of reads on a B200, printing the PAF lines of src/rmap.cpp:954-972.
//
//   rawalign_b200_map [options] <ref.ind> <reads.bin>
//
// <ref.ind> is the reference's unchanged index file (rawalign -d).  <reads.bin> holds already-extracted DAC samples (the
// FAST5/HDF5 reader stays the reference's host code and is not part of this path):
//   u32 n_reads | u64 raw_off[n] | u32 raw_len[n] | f32 offset[n] | f32 scale[n] | u64 n_raw | i16 raw[n_raw]
// Options mirror the reference's flags (src/main.cpp:231-251): --dtw-evaluate-chains, --dtw-border-constraint global|sparse,
// --dtw-fill-method full|banded=FRAC, --dtw-match-bonus F, --dtw-min-score F, --max-chunks N, --min-events N, --batch N.
// Tags are the reference's (rmap.cpp:730-758) except mt:f: (per-read wall time has no meaning in a batch; the batch time per
// read is printed instead, a documented deviation).
#include "rawalign_b200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static int die(const char *what) { fprintf(stderr, "rawalign_b200_map: %s: %s\n", what, ra_last_error()); return 1; }

// sequence names and lengths from the index header (rawindex.cpp:283-296)
static bool index_names(const char *path, std::vector<std::string> &names, std::vector<uint32_t> &lens)
{
	FILE *f = fopen(path, "rb");
	if (!f) return false;
	char magic[2]; uint32_t pars[8];
	if (fread(magic, 1, 2, f) != 2 || fread(pars, 4, 8, f) != 8) { fclose(f); return false; }
	for (uint32_t i = 0; i < pars[6]; ++i) {
		uint8_t l; uint32_t len;
		if (fread(&l, 1, 1, f) != 1) { fclose(f); return false; }
		std::string nm(l, ' ');
		if (l && fread(&nm[0], 1, l, f) != l) { fclose(f); return false; }
		if (fread(&len, 4, 1, f) != 1) { fclose(f); return false; }
		fseek(f, 8L * len, SEEK_CUR);
		names.push_back(nm); lens.push_back(len);
	}
	fclose(f);
	return true;
}

int main(int argc, char **argv)
{
	ra_ri_mapopt_t opt; ra_ri_mapopt_init(&opt);
	uint32_t batch = 8192;
	int a = 1;
	for (; a < argc && argv[a][0] == '-'; ++a) {
		std::string o = argv[a];
		auto val = [&]() -> const char * { return a + 1 < argc ? argv[++a] : ""; };
		if (o == "--dtw-evaluate-chains") opt.flag |= RA_M_DTW_EVALUATE_CHAINS;
		else if (o == "--dtw-border-constraint") { std::string v = val(); opt.dtw_border_constraint = v == "global" ? RA_DTW_BORDER_CONSTRAINT_GLOBAL : RA_DTW_BORDER_CONSTRAINT_SPARSE; }
		else if (o == "--dtw-fill-method") {
			std::string v = val();
			if (v == "full") opt.dtw_fill_method = RA_DTW_FILL_METHOD_FULL;
			else if (v.rfind("banded", 0) == 0) { opt.dtw_fill_method = RA_DTW_FILL_METHOD_BANDED; if (v.size() > 7) opt.dtw_band_radius_frac = (float)atof(v.c_str() + 7); }
		}
		else if (o == "--dtw-match-bonus") opt.dtw_match_bonus = (float)atof(val());
		else if (o == "--dtw-min-score") opt.dtw_min_score = (float)atof(val());
		else if (o == "--max-chunks") opt.max_num_chunk = (uint32_t)atoi(val());
		else if (o == "--min-events") opt.min_events = (uint32_t)atoi(val());
		else if (o == "--batch") batch = (uint32_t)atoi(val());
		else { fprintf(stderr, "unknown option %s\n", o.c_str()); return 2; }
	}
	if (argc - a < 2) { fprintf(stderr, "usage: rawalign_b200_map [options] <ref.ind> <reads.bin>\n"); return 2; }
	const char *ind = argv[a], *reads = argv[a + 1];
	std::vector<std::string> names; std::vector<uint32_t> lens;
	if (!index_names(ind, names, lens)) { fprintf(stderr, "cannot read %s\n", ind); return 1; }

	FILE *f = fopen(reads, "rb");
	if (!f) { fprintf(stderr, "cannot open %s\n", reads); return 1; }
	uint32_t n = 0; uint64_t n_raw = 0;
	if (fread(&n, 4, 1, f) != 1) return 1;
	std::vector<uint64_t> off(n); std::vector<uint32_t> len(n); std::vector<float> dac_off(n), dac_scale(n);
	if (fread(off.data(), 8, n, f) != n || fread(len.data(), 4, n, f) != n || fread(dac_off.data(), 4, n, f) != n || fread(dac_scale.data(), 4, n, f) != n) return 1;
	if (fread(&n_raw, 8, 1, f) != 1) return 1;
	std::vector<int16_t> raw(n_raw);
	if (fread(raw.data(), 2, n_raw, f) != n_raw) return 1;
	fclose(f);

	ra_ctx *ctx = nullptr; ra_index *ix = nullptr;
	if (ra_init(0, &ctx)) return die("ra_init");
	if (ra_index_load_file(ctx, ind, &ix)) return die("ra_index_load_file");
	std::vector<ra_paf_t> rec(n);
	double total_ms = 0;
	for (uint32_t r0 = 0; r0 < n; r0 += batch) {                      // mini-batches, as ri_sig_read_frag cuts them (rmap.cpp:824)
		const uint32_t m = std::min(batch, n - r0);
		ra_map_stats_t st;
		if (ra_map_batch(ctx, ix, &opt, raw.data(), off.data() + r0, len.data() + r0, dac_off.data() + r0, dac_scale.data() + r0,
		                 nullptr, nullptr, nullptr, m, rec.data() + r0, &st)) return die("ra_map_batch");
		total_ms += st.total_ms;
	}
	const double mt = n ? total_ms / n : 0.0;
	for (uint32_t r = 0; r < n; ++r) {                                  // rmap.cpp:954-972 with the tags of :730-758 / :764-790
		const ra_paf_t &o = rec[r];
		char name[32]; snprintf(name, sizeof(name), "read_%u", r);
		if (o.mapped)
			printf("%s\t%u\t%u\t%u\t%c\t%s\t%u\t%u\t%u\t%u\t%u\t%d\tmt:f:%f\tci:i:%u\tsl:i:%u\tcm:i:%u\tnc:i:%u\ts1:f:%f\ts2:f:%f\tsm:f:%f\tat:f:%f\taq:f:%f\n",
			       name, o.read_length, o.read_start_position, o.read_end_position, o.rev ? '-' : '+', names[o.ref_id].c_str(), lens[o.ref_id],
			       o.fragment_start_position, o.fragment_start_position + o.fragment_length, o.read_end_position - o.read_start_position - 1,
			       o.fragment_length, (int)o.mapq, mt, o.ci, o.sl, o.cm, o.nc, o.s1, o.s2, o.sm, o.at, o.aq);
		else if (o.nc >= 1)
			printf("%s\t%u\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\tmt:f:%f\tci:i:%u\tsl:i:%u\tcm:i:%u\tnc:i:%u\ts1:f:%f\ts2:f:%f\tsm:f:%f\tat:f:%f\taq:f:%f\n",
			       name, o.read_length, (int)o.mapq, mt, o.ci, o.sl, o.cm, o.nc, o.s1, o.s2, o.sm, o.at, o.aq);
		else
			printf("%s\t%u\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\tmt:f:%f\tci:i:%u\tsl:i:%u\tcm:i:0\tnc:i:0\ts1:f:0\ts2:f:0\tsm:f:0\tat:f:0\taq:f:0\n",
			       name, o.read_length, (int)o.mapq, mt, o.ci, o.sl);
	}
	ra_index_destroy(ix);
	ra_destroy(ctx);
	return 0;
}
