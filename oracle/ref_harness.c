/*
 * ref_harness.c -- thin driver around the UNMODIFIED reference objects (TEST INFRASTRUCTURE ONLY).
 *
 * oracle/Makefile compiles /root/reference/{chain.c,kalloc.c,misc.c,kthread.c} where they lie and links them
 * with this file into oracle/_ref/libmm2ref_chain.so.  No reference source is copied into this repository.
 * The harness only (a) supplies the three globals kthread.c expects (kthread.c:11-13), (b) makes the
 * consuming entry points callable on caller-owned arrays, and (c) runs the reference mm_chain_dp over a batch
 * of reads under the reference's own kt_for (kthread.c:125) for the timed CPU baseline.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* declarations copied by meaning, not by text: minimap.h:48-55, mmpriv.h:68-70, kalloc.h, kthread.h, misc.c:136-140 */
typedef struct { uint64_t x, y; } mm128_t;
struct new_seed { mm128_t seed; int32_t p; int32_t f; };
mm128_t *mm_chain_dp(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_cnt, int min_sc, int is_cdna,
                     int n_segs, int64_t n, mm128_t *a, int *n_u_, uint64_t **_u, void *km);
struct new_seed *mm_chain_dp_fpga(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_sc, int is_cdna,
                                  int n_segs, int64_t n, mm128_t *a, uint32_t *_new_i);
mm128_t *mm_chain_dp_bottom(int min_cnt, int min_sc, int n_segs, int *n_u_, uint64_t **_u, void *km,
                            struct new_seed *fpga_a, uint32_t new_i);
void *kmalloc(void *km, size_t size);
void kfree(void *km, void *ptr);
void *km_init(void);
void km_destroy(void *km);
void kt_for(int n_threads, void (*func)(void*, long, int), void *data, long n);

double send_task1[100], send_task2[100], process_result[100];   /* kthread.c:11-13 expects these */

typedef struct {
	int32_t max_dist_x, max_dist_y, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs, max_iter;
} ref_params_t;   /* same layout as orc_params_t; max_iter must be 0 (the reference has no such clamp) */

static double now_s(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static int64_t chain_len_sum(int n_u, const uint64_t *u)
{
	int64_t s = 0; int i;
	for (i = 0; i < n_u; ++i) s += (int32_t)u[i];
	return s;
}

/* one call of the monolithic reference function on a private copy of a[] (it frees its input, chain.c:137,214) */
int32_t ref_chain_dp(const ref_params_t *prm, int64_t n, const mm128_t *a, uint64_t **u, mm128_t **b, int64_t *n_b)
{
	int n_u = 0;
	mm128_t *copy = (mm128_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(mm128_t));
	if (n > 0) memcpy(copy, a, (size_t)n * sizeof(mm128_t));
	*u = 0;
	*b = mm_chain_dp(prm->max_dist_x, prm->max_dist_y, prm->bw, prm->max_skip, prm->min_cnt, prm->min_sc,
	                 prm->is_cdna, prm->n_segs, n, copy, &n_u, u, 0);
	*n_b = chain_len_sum(n_u, *u);
	return n_u;
}

/* the fork's split form: device half */
uint32_t ref_chain_dp_fpga(const ref_params_t *prm, int64_t n, const mm128_t *a, struct new_seed **rec)
{
	uint32_t new_i = 0;
	mm128_t *copy = (mm128_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(mm128_t));
	if (n > 0) memcpy(copy, a, (size_t)n * sizeof(mm128_t));
	*rec = mm_chain_dp_fpga(prm->max_dist_x, prm->max_dist_y, prm->bw, prm->max_skip, prm->min_sc,
	                        prm->is_cdna, prm->n_segs, n, copy, &new_i);
	return new_i;
}

/* the fork's split form: host half (does not free rec, chain.c:408) */
int32_t ref_chain_dp_bottom(const ref_params_t *prm, uint32_t new_i, struct new_seed *rec, uint64_t **u, mm128_t **b, int64_t *n_b)
{
	int n_u = 0;
	*u = 0;
	*b = mm_chain_dp_bottom(prm->min_cnt, prm->min_sc, prm->n_segs, &n_u, u, 0, rec, new_i);
	*n_b = chain_len_sum(n_u, *u);
	return n_u;
}

void ref_free(void *p) { free(p); }

/* ---- batch under the reference's kt_for ------------------------------------------------------------------ */

typedef struct {
	const ref_params_t *prm;
	const int64_t *offs;
	const mm128_t *a;
	mm128_t **copies;      /* pre-made malloc copies (km == NULL mode) or NULL (arena mode: copy in the worker) */
	void **km;             /* per-thread arenas (arena mode) or NULL */
	mm128_t **b; uint64_t **u; int32_t *n_u; int32_t *tid;
} batch_t;

static void batch_worker(void *data, long r, int tid)
{
	batch_t *bt = (batch_t*)data;
	const ref_params_t *prm = bt->prm;
	int64_t n = bt->offs[r+1] - bt->offs[r];
	void *km = bt->km ? bt->km[tid] : 0;
	mm128_t *in;
	int n_u = 0;
	if (bt->copies) in = bt->copies[r];
	else {
		in = (mm128_t*)kmalloc(km, (size_t)(n > 0 ? n : 1) * sizeof(mm128_t));
		if (n > 0) memcpy(in, bt->a + bt->offs[r], (size_t)n * sizeof(mm128_t));
	}
	bt->u[r] = 0;
	bt->b[r] = mm_chain_dp(prm->max_dist_x, prm->max_dist_y, prm->bw, prm->max_skip, prm->min_cnt, prm->min_sc,
	                       prm->is_cdna, prm->n_segs, n, in, &n_u, &bt->u[r], km);
	bt->n_u[r] = n_u;
	bt->tid[r] = tid;
}

/* Runs the reference over all reads with n_threads kt_for workers.  use_kalloc=1: one km_init() arena per
 * thread, input copied into the arena inside the timed region (what map.c does: anchors are built in the
 * arena); use_kalloc=0: km=NULL (the authors' --no-kalloc run, run.sh:3), inputs pre-copied outside the timed
 * region.  Packed outputs like orc_chain_batch; u_out/b_out malloc()ed (pass NULL to discard).
 * Returns the wall time in seconds of the kt_for region only. */
double ref_chain_batch_mt(const ref_params_t *prm, int64_t n_reads, const int64_t *offs, const mm128_t *a,
                          int n_threads, int use_kalloc,
                          int32_t *n_u, int64_t *u_off, int64_t *b_off, uint64_t **u_out, mm128_t **b_out)
{
	batch_t bt;
	int64_t r, tu = 0, tb = 0;
	int t;
	double t0, t1;
	if (n_threads < 1) n_threads = 1;
	memset(&bt, 0, sizeof(bt));
	bt.prm = prm, bt.offs = offs, bt.a = a;
	bt.b = (mm128_t**)calloc((size_t)n_reads + 1, sizeof(void*));
	bt.u = (uint64_t**)calloc((size_t)n_reads + 1, sizeof(void*));
	bt.n_u = (int32_t*)calloc((size_t)n_reads + 1, sizeof(int32_t));
	bt.tid = (int32_t*)calloc((size_t)n_reads + 1, sizeof(int32_t));
	if (use_kalloc) {
		bt.km = (void**)calloc((size_t)n_threads, sizeof(void*));
		for (t = 0; t < n_threads; ++t) bt.km[t] = km_init();
	} else {
		bt.copies = (mm128_t**)calloc((size_t)n_reads + 1, sizeof(void*));
		for (r = 0; r < n_reads; ++r) {
			int64_t n = offs[r+1] - offs[r];
			bt.copies[r] = (mm128_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(mm128_t));
			if (n > 0) memcpy(bt.copies[r], a + offs[r], (size_t)n * sizeof(mm128_t));
		}
	}
	t0 = now_s();
	kt_for(n_threads, batch_worker, &bt, (long)n_reads);
	t1 = now_s();

	for (r = 0; r < n_reads; ++r) {
		if (u_off) u_off[r] = tu;
		if (b_off) b_off[r] = tb;
		tu += bt.n_u[r], tb += chain_len_sum(bt.n_u[r], bt.u[r]);
		if (n_u) n_u[r] = bt.n_u[r];
	}
	if (u_off) u_off[n_reads] = tu;
	if (b_off) b_off[n_reads] = tb;
	if (u_out && b_out) {
		uint64_t *U = (uint64_t*)malloc((size_t)(tu > 0 ? tu : 1) * 8);
		mm128_t *B = (mm128_t*)malloc((size_t)(tb > 0 ? tb : 1) * sizeof(mm128_t));
		for (r = 0, tu = tb = 0; r < n_reads; ++r) {
			int64_t nb = chain_len_sum(bt.n_u[r], bt.u[r]);
			if (bt.n_u[r]) memcpy(U + tu, bt.u[r], (size_t)bt.n_u[r] * 8);
			if (nb) memcpy(B + tb, bt.b[r], (size_t)nb * sizeof(mm128_t));
			tu += bt.n_u[r], tb += nb;
		}
		*u_out = U, *b_out = B;
	}
	for (r = 0; r < n_reads; ++r) {
		void *km = bt.km ? bt.km[bt.tid[r]] : 0;
		kfree(km, bt.b[r]); kfree(km, bt.u[r]);
	}
	if (bt.km) { for (t = 0; t < n_threads; ++t) km_destroy(bt.km[t]); free(bt.km); }
	free(bt.copies); free(bt.b); free(bt.u); free(bt.n_u); free(bt.tid);
	return t1 - t0;
}
