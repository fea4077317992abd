// ThreadSanitizer stress of the library's host side (tests/test_tsan_host.py builds and runs it; no GPU involved).
//   * T threads hammer the three drop-in symbols at once (mm_chain_dp from "workers", mm_chain_dp_fpga + mm_chain_dp_bottom
//     from others, two parameter sets) -> the combining dispatchers, their private contexts, per-call result hand-over;
//   * meanwhile other threads run mm2b200_chain_batch on their own contexts with tiny sub-batches (slot pipeline) and
//     mm2b200_chain_batch_multi over several contexts (host thread per context).
// Every call checks that it got its own read back.  Exit code 0 and no ThreadSanitizer report = pass.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include "mm2chain_b200.h"

static std::atomic<int> bad{0};

static void fill_read(mm128_t *a, int n, uint64_t tag)
{
	for (int i = 0; i < n; ++i) { a[i].x = (tag << 32) | (uint64_t)(100 + i); a[i].y = ((uint64_t)15 << 32) | (uint64_t)(10 + i); }
}

static void entry_worker(int t, int iters)
{
	for (int it = 0; it < iters; ++it) {
		const uint64_t tag = (uint64_t)t * 100000 + (uint64_t)it + 1;
		const int n = 3 + (int)(tag % 50);
		mm128_t *a = (mm128_t*)malloc(sizeof(mm128_t) * n);
		fill_read(a, n, tag);
		const uint64_t x0 = a[0].x;
		if (t % 3 == 0) {
			int n_u = 0; uint64_t *u = 0;
			mm128_t *b = mm_chain_dp(5000, 5000, 500, 25, (t & 4) ? 1 : 3, 40, 0, 1, n, a, &n_u, &u, 0);
			if (!b || n_u != 1 || b[0].x != x0 || u[0] != (((uint64_t)1 << 32) | 1)) ++bad;
			free(b); free(u);
		} else {
			uint32_t new_i = 0;
			struct new_seed *rec = mm_chain_dp_fpga(5000, 5000, 500, 25, 40, 0, 1, n, a, &new_i);
			if (!rec || (int)new_i != n || rec[0].seed.x != x0 || rec[n - 1].f != n - 1) ++bad;
			int n_u = 0; uint64_t *u = 0;
			mm128_t *b = rec ? mm_chain_dp_bottom(3, 40, 1, &n_u, &u, 0, rec, new_i) : 0;
			if (!b || n_u != 1 || b[0].x != x0) ++bad;
			free(b); free(u); free(rec);
		}
	}
}

static void batch_worker(int t, int iters)
{
	mm2b200_config_t cfg; memset(&cfg, 0, sizeof(cfg));
	cfg.n_slots = 2 + t % 2; cfg.slot_anchors = 64;      // tiny sub-batches: the slot pipeline wraps around many times
	mm2b200_ctx_t *ctx = 0;
	if (mm2b200_create(&cfg, &ctx)) { ++bad; return; }
	mm2b200_params_t prm = {5000, 5000, 500, 25, 3, 40, 0, 1, 0};
	for (int it = 0; it < iters; ++it) {
		const int R = 40;
		std::vector<int64_t> offs(R + 1, 0);
		for (int r = 0; r < R; ++r) offs[r + 1] = offs[r] + 1 + (r * 7 + it) % 30;
		std::vector<mm128_t> a((size_t)offs[R]);
		for (int r = 0; r < R; ++r) fill_read(a.data() + offs[r], (int)(offs[r + 1] - offs[r]), (uint64_t)(t * 1000 + r + 1));
		mm2b200_result_t res;
		if (mm2b200_chain_batch(ctx, &prm, R, offs.data(), a.data(), &res)) { ++bad; continue; }
		for (int r = 0; r < R; ++r)
			if (res.n_u[r] != 1 || res.b[res.b_off[r]].x != a[offs[r]].x) ++bad;
	}
	mm2b200_destroy(ctx);
}

static void multi_worker(int t, int iters)
{
	const int N = 3;
	mm2b200_ctx_t *ctxs[N];
	mm2b200_config_t cfg; memset(&cfg, 0, sizeof(cfg));
	cfg.n_slots = 2; cfg.slot_anchors = 128;
	for (int k = 0; k < N; ++k) { cfg.device = k; if (mm2b200_create(&cfg, &ctxs[k])) { ++bad; return; } }
	mm2b200_params_t prm = {5000, 5000, 500, 25, 3, 40, 0, 1, 0};
	for (int it = 0; it < iters; ++it) {
		const int R = 90;
		std::vector<int64_t> offs(R + 1, 0);
		for (int r = 0; r < R; ++r) offs[r + 1] = offs[r] + 1 + (r * 5 + it + t) % 20;
		std::vector<mm128_t> a((size_t)offs[R]);
		for (int r = 0; r < R; ++r) fill_read(a.data() + offs[r], (int)(offs[r + 1] - offs[r]), (uint64_t)(t * 1000 + r + 1));
		int64_t cuts[N + 1];
		mm2b200_result_t res[N];
		if (mm2b200_chain_batch_multi(ctxs, N, &prm, R, offs.data(), a.data(), cuts, res)) { ++bad; continue; }
		for (int k = 0; k < N; ++k)
			for (int64_t r = cuts[k]; r < cuts[k + 1]; ++r)
				if (res[k].n_u[r - cuts[k]] != 1 || res[k].b[res[k].b_off[r - cuts[k]]].x != a[offs[r]].x) ++bad;
	}
	for (int k = 0; k < N; ++k) mm2b200_destroy(ctxs[k]);
}

int main(int argc, char **argv)
{
	const int T = argc > 1 ? atoi(argv[1]) : 24, iters = argc > 2 ? atoi(argv[2]) : 200;
	std::vector<std::thread> th;
	for (int t = 0; t < T; ++t) th.emplace_back(entry_worker, t, iters);
	for (int t = 0; t < 4; ++t) th.emplace_back(batch_worker, t, iters / 4 + 1);
	for (int t = 0; t < 2; ++t) th.emplace_back(multi_worker, t, iters / 8 + 1);
	for (auto &x : th) x.join();
	int64_t nb = 0, nc = 0;
	mm2b200_entry_stats(&nb, &nc);
	printf("stress: %d mismatches; drop-in calls %lld in %lld device batches\n", bad.load(), (long long)nc, (long long)nb);
	return bad.load() ? 1 : 0;
}
