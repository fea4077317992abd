/* dropin_rate.c -- level-1 drop-in throughput from C: T pthreads each call the reference entry points one read at a
 * time, exactly like the worker / result threads of map.c do: mm_chain_dp (chain.c:23; "chain"), or the fork's live pair
 * mm_chain_dp_fpga + mm_chain_dp_bottom (map.c:525,837; "split").
 *   gcc -O2 -I include tools/dropin_rate.c -o /tmp/dropin_rate -L minimap2_chaindp_b200 -lmm2chain_b200 \
 *       -Wl,-rpath,$PWD/minimap2_chaindp_b200 -lpthread && /tmp/dropin_rate anchors.bin [chain|split]
 * anchors.bin: int64 n_reads, int64 offs[n_reads+1], mm128_t a[offs[n_reads]] (written by tools/dropin_rate.py --dump). */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "mm2chain_b200.h"

static int64_t n_reads, *offs;
static mm128_t *a;
static int T, split_mode;
static long chains[1024];

static void *worker(void *arg)
{
	long t = (long)arg, r, nc = 0;
	for (r = t; r < n_reads; r += T) {
		int64_t n = offs[r + 1] - offs[r];
		mm128_t *in = (mm128_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(mm128_t)), *b;
		uint64_t *u = 0;
		int n_u = 0;
		memcpy(in, a + offs[r], (size_t)n * sizeof(mm128_t));
		if (split_mode) {
			uint32_t new_i = 0;
			struct new_seed *rec = mm_chain_dp_fpga(5000, 5000, 500, 25, 40, 0, 1, n, in, &new_i);   /* consumes in */
			b = mm_chain_dp_bottom(3, 40, 1, &n_u, &u, 0, rec, new_i);
			free(rec);
		} else b = mm_chain_dp(5000, 5000, 500, 25, 3, 40, 0, 1, n, in, &n_u, &u, 0);   /* consumes in */
		nc += n_u;
		free(b); free(u);
	}
	chains[t] = nc;
	return 0;
}

int main(int argc, char **argv)
{
	FILE *fp = fopen(argc > 1 ? argv[1] : "anchors.bin", "rb");
	int ts[] = {1, 8, 64, 128, 512}, k;
	split_mode = argc > 2 && !strcmp(argv[2], "split");
	if (!fp || fread(&n_reads, 8, 1, fp) != 1) return 1;
	offs = (int64_t*)malloc((size_t)(n_reads + 1) * 8);
	if (fread(offs, 8, (size_t)n_reads + 1, fp) != (size_t)n_reads + 1) return 1;
	a = (mm128_t*)malloc((size_t)offs[n_reads] * sizeof(mm128_t));
	if (fread(a, sizeof(mm128_t), (size_t)offs[n_reads], fp) != (size_t)offs[n_reads]) return 1;
	for (k = 0; k < 5; ++k) {
		pthread_t th[512];
		struct timespec t0, t1;
		long i, total = 0;
		double dt;
		T = ts[k];
		clock_gettime(CLOCK_MONOTONIC, &t0);
		for (i = 0; i < T; ++i) pthread_create(&th[i], 0, worker, (void*)i);
		for (i = 0; i < T; ++i) pthread_join(th[i], 0);
		clock_gettime(CLOCK_MONOTONIC, &t1);
		for (i = 0; i < T; ++i) total += chains[i];
		dt = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
		printf("%4d caller threads: %9.0f reads/s through %s (%ld chains%s)\n", T, n_reads / dt,
		       split_mode ? "mm_chain_dp_fpga + mm_chain_dp_bottom" : "mm_chain_dp", total, mm2b200_last_error()[0] ? ", ERROR" : "");
	}
	return 0;
}
