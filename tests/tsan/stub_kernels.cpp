// Stand-ins for the launchers of csrc/chain_kernels.cu (tests/test_tsan_host.py only; see cuda_runtime.h next to this
// file).  They do NOT chain anything: each read with at least one anchor "yields" one chain made of its first anchor
// (u = 1<<32 | 1), each read's record list is its anchors with f = p = position.  That is enough for every caller to check
// that the host machinery handed it ITS read's data, which is what the race detector run is about.
#include "chain_internal.h"

namespace mm2b200 {

int launch_prep(const Workspace &w, const mm2b200_params_t &, cudaStream_t) { w.counters[CNT_UNITS] = w.n_reads; return 1; }
int launch_fill(const Workspace &, const mm2b200_params_t &, const LaunchCfg &, bool, cudaStream_t) { return 1; }

int launch_compact(const Workspace &w, const mm2b200_params_t &, cudaStream_t)
{
	for (int r = 0; r < w.n_reads; ++r) {
		const int64_t b = w.offs[r], e = w.offs[r + 1];
		for (int64_t i = b; i < e; ++i) {
			new_seed &s = w.rec[i];
			s.seed.x = w.a[i].x; s.seed.y = w.a[i].y; s.p = (int32_t)(i - b); s.f = (int32_t)(i - b);
		}
		w.new_i[r] = (uint32_t)(e - b);
	}
	return 1;
}

int launch_split_records(const Workspace &w, cudaStream_t)
{
	for (int r = 0; r < w.n_reads; ++r) w.new_i[r] = (uint32_t)(w.offs[r + 1] - w.offs[r]);
	return 1;
}

int launch_tail(const Workspace &w, const mm2b200_params_t &, cudaStream_t)
{
	long long tu = 0;
	for (int r = 0; r < w.n_reads; ++r) {
		const int n = w.new_i[r] > 0 ? 1 : 0;
		w.n_u[r] = n; w.n_v[r] = n; w.u_off[r] = tu; w.b_off[r] = tu; tu += n;
	}
	w.u_off[w.n_reads] = tu; w.b_off[w.n_reads] = tu;
	w.totals[0] = tu; w.totals[1] = tu; w.totals[2] = w.n_reads; w.totals[3] = 0; w.totals[4] = 0; w.totals[5] = 0;
	return 1;
}

int launch_tail_reads(const Workspace &w, const mm2b200_params_t &p, cudaStream_t s) { return launch_tail(w, p, s); }
int launch_tail_scan(const Workspace &, cudaStream_t) { return 1; }

int launch_gather(const Workspace &w, cudaStream_t)
{
	for (int r = 0; r < w.n_reads; ++r)
		if (w.n_u[r]) {
			const new_seed &s = w.rec[w.offs[r]];
			w.u_out[w.u_off[r]] = ((uint64_t)1 << 32) | 1;
			w.b_out[w.b_off[r]].x = s.seed.x; w.b_out[w.b_off[r]].y = s.seed.y;
		}
	return 1;
}

int launch_pack_records(const Workspace &w, cudaStream_t)
{
	long long acc = 0;
	for (int r = 0; r < w.n_reads; ++r) {
		w.rec_off[r] = acc;
		for (uint32_t k = 0; k < w.new_i[r]; ++k) w.rec_out[acc + k] = w.rec[w.offs[r] + k];
		acc += w.new_i[r];
	}
	w.rec_off[w.n_reads] = acc;
	w.totals[0] = 0; w.totals[1] = 0; w.totals[2] = w.n_reads; w.totals[3] = acc; w.totals[4] = 0; w.totals[5] = 0;
	return 1;
}

int launch_export_fp(const Workspace &, cudaStream_t) { return 1; }
int launch_unpack_wire(const Workspace &, const void *, size_t, size_t, cudaStream_t) { return 1; }
void query_launch_cfg(LaunchCfg *cfg) { *cfg = LaunchCfg(); cfg->sm_count = 1; cfg->fill_blocks_per_sm = 1; }
long long launch_int_peak(int, int, int, unsigned *, cudaStream_t) { return 1; }

} // namespace mm2b200
