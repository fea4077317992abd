/*
 * chain_oracle.c -- CPU restatement of minimap2's anchor-chaining DP.
 *
 * TEST INFRASTRUCTURE ONLY (see chain_oracle.h).  Every function cites the reference lines it restates
 * (paths are into /root/reference).  The code is written from the specification in SURVEY.md Appendix A/B,
 * organised as separate stages (pair scoring, predecessor scan state machine, compaction, tail) so that the
 * CUDA kernels can be compared stage by stage.
 *
 * Build: see oracle/Makefile (plain gcc -O2; no -ffast-math, no FMA contraction matters here because the
 * only floating-point expression is a chain of two multiplies, chain.c:73,78).
 */
#include <stdlib.h>
#include <string.h>
#include "chain_oracle.h"

#define SEG_OF(y)   ((int32_t)(((y) >> 48) & 0xff))   /* MM_SEED_SEG_MASK/SHIFT, mmpriv.h:21-22 */
#define SPAN_OF(y)  ((int32_t)(((y) >> 32) & 0xff))   /* chain.c:55 "only 8 bits of span is used" */
#define QPOS_OF(y)  ((int32_t)(y))

/* floor(log2(v)), v >= 1 -- same value as ilog2_32 (chain.c:16-21), computed without the byte table */
static inline int floor_log2_u32(uint32_t v)
{
	int r = 0;
	while (v >>= 1) ++r;
	return r;
}

float orc_avg_qspan(int64_t n, const orc_anchor_t *a)
{	/* chain.c:46-47: u64 sum -> float, int64 n -> float, float divide */
	uint64_t sum = 0;
	int64_t i;
	for (i = 0; i < n; ++i) sum += (uint64_t)SPAN_OF(a[i].y);
	return (float)sum / n;
}

/* Outcome of looking at predecessor j from anchor i, before f[j] is added (order-free part).
 * Returns 0 if the pair is skipped by one of the `continue`s at chain.c:63-67, else 1 with *base set to
 * the score contribution of chain.c:68-78. */
static inline int pair_base(const orc_params_t *prm, float avg_qspan, const orc_anchor_t *ai, const orc_anchor_t *aj, int32_t *base)
{
	int64_t dr = (int64_t)(ai->x - aj->x);                       /* chain.c:60 */
	int32_t dq = (int32_t)((uint32_t)QPOS_OF(ai->y) - (uint32_t)QPOS_OF(aj->y)); /* chain.c:61 */
	int32_t span_i = SPAN_OF(ai->y);
	int same_seg = SEG_OF(ai->y) == SEG_OF(aj->y);
	int32_t dd, sc, log_dd, lin;
	if ((same_seg && dr == 0) || dq <= 0) return 0;               /* chain.c:63 */
	if ((same_seg && dq > prm->max_dist_y) || dq > prm->max_dist_x) return 0; /* chain.c:64 */
	dd = (int32_t)(dr > dq ? dr - dq : dq - dr);                  /* chain.c:65 */
	if (same_seg && dd > prm->bw) return 0;                       /* chain.c:66 */
	if (prm->n_segs > 1 && !prm->is_cdna && same_seg && dr > prm->max_dist_y) return 0; /* chain.c:67 */
	sc = (int32_t)(dq < dr ? dq : dr);                            /* chain.c:68-69 */
	if (sc > span_i) sc = span_i;
	log_dd = dd ? floor_log2_u32((uint32_t)dd) : 0;               /* chain.c:70 */
	lin = (int32_t)(dd * .01 * avg_qspan);                        /* chain.c:73,78: ((double)dd*0.01)*(double)avg */
	if (prm->is_cdna || !same_seg) {                              /* chain.c:71-77 */
		if (!same_seg && dr == 0) ++sc;
		else if (dr > dq || !same_seg) sc -= lin < log_dd ? lin : log_dd;
		else sc -= lin + (log_dd >> 1);
	} else sc -= lin + (log_dd >> 1);                             /* chain.c:78 */
	*base = sc;
	return 1;
}

int64_t orc_fill(const orc_params_t *prm, int64_t n, const orc_anchor_t *a,
                 int32_t *f, int32_t *p, int32_t *v, int32_t *visited, int32_t *window)
{
	int64_t i, j, st = 0, total = 0;
	int32_t *mark;
	float avg_qspan;
	if (n <= 0) return 0;
	avg_qspan = orc_avg_qspan(n, a);
	mark = (int32_t*)malloc((size_t)n * sizeof(int32_t));
	for (i = 0; i < n; ++i) mark[i] = -1;   /* reference zero-fills and tags with i (chain.c:40,83,87); i=0 has an
	                                           empty window, so any never-matching initial tag is equivalent */
	for (i = 0; i < n; ++i) {
		int32_t span_i = SPAN_OF(a[i].y);
		int32_t best = span_i, n_skip = 0, n_vis = 0;
		int64_t best_j = -1;
		while (st < i && a[i].x - a[st].x > (uint64_t)(int64_t)prm->max_dist_x) ++st;   /* chain.c:58 (unsigned compare) */
		if (prm->max_iter > 0 && i - st > prm->max_iter) st = i - prm->max_iter;        /* extension; off when 0 */
		if (window) window[i] = (int32_t)(i - st);
		for (j = i - 1; j >= st; --j) {                                                  /* chain.c:59 */
			int32_t base, sc;
			++n_vis;
			if (!pair_base(prm, avg_qspan, &a[i], &a[j], &base)) continue;              /* SKIP: no side effects */
			sc = base + f[j];                                                            /* chain.c:79 */
			if (sc > best) {                                                             /* chain.c:80-82 IMPROVE */
				best = sc, best_j = j;
				if (n_skip > 0) --n_skip;
			} else if (mark[j] == (int32_t)i) {                                          /* chain.c:83-86 MARKED-MISS */
				if (++n_skip > prm->max_skip) break;
			}
			if (p[j] >= 0) mark[p[j]] = (int32_t)i;                                      /* chain.c:87 */
		}
		f[i] = best, p[i] = (int32_t)best_j;                                             /* chain.c:89 */
		v[i] = (best_j >= 0 && v[best_j] > best) ? v[best_j] : best;                     /* chain.c:90 */
		if (visited) visited[i] = n_vis;
		total += n_vis;
	}
	free(mark);
	return total;
}

uint32_t orc_compact(int32_t min_sc, int64_t n, const orc_anchor_t *a,
                     const int32_t *f, const int32_t *p, const int32_t *v,
                     orc_seed_t *rec, int32_t *newid_out)
{	/* chain.c:93-122.  Emission order defines the new indices: a not-yet-emitted predecessor first, then i. */
	int64_t i;
	uint32_t m = 0;
	int32_t *newid = newid_out ? newid_out : (int32_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int32_t));
	for (i = 0; i < n; ++i) newid[i] = -1;
	for (i = 0; i < n; ++i) {
		int32_t k = p[i];
		if (k >= 0 && newid[k] < 0) {                    /* chain.c:97-107: late predecessor */
			rec[m].seed = a[k];
			rec[m].f = f[k];
			rec[m].p = (int32_t)((uint32_t)-1 << 2) | (v[k] >= min_sc) | ((f[k] < v[k]) << 1);
			newid[k] = (int32_t)m++;
		}
		if (v[i] >= min_sc || k >= 0) {                  /* chain.c:109-122 */
			rec[m].seed = a[i];
			rec[m].f = f[i];
			rec[m].p = (int32_t)((uint32_t)(k >= 0 ? newid[k] : -1) << 2) | (v[i] >= min_sc) | ((f[i] < v[i]) << 1);
			newid[i] = (int32_t)m++;
		}
	}
	if (!newid_out) free(newid);
	return m;
}

/* ---- ksort.h:101-151 restated with indices instead of pointers ------------------------------------------- */

#define ORC_RS_MIN 64   /* RS_MIN_SIZE, ksort.h:98 */

#define DEFINE_RADIX(NAME, TYPE, KEY)                                                                    \
static void ins_##NAME(TYPE *z, int64_t n)                                                               \
{	/* rs_insertsort, ksort.h:105-115 (stable) */                                                       \
	int64_t i, j;                                                                                        \
	for (i = 1; i < n; ++i) {                                                                            \
		if (KEY(z[i]) < KEY(z[i-1])) {                                                                   \
			TYPE tmp = z[i];                                                                             \
			for (j = i; j > 0 && KEY(tmp) < KEY(z[j-1]); --j) z[j] = z[j-1];                             \
			z[j] = tmp;                                                                                  \
		}                                                                                                \
	}                                                                                                    \
}                                                                                                        \
static void msd_##NAME(TYPE *z, int64_t n, int shift)                                                    \
{	/* rs_sort, ksort.h:116-146: one 8-bit American-flag pass in place, then recurse / insertion sort */ \
	int64_t head[256], tail[256], lo[257];                                                               \
	int k;                                                                                               \
	int64_t i;                                                                                           \
	memset(lo, 0, sizeof(lo));                                                                           \
	for (i = 0; i < n; ++i) ++lo[((KEY(z[i]) >> shift) & 0xff) + 1];                                     \
	for (k = 0; k < 256; ++k) lo[k+1] += lo[k];                                                          \
	for (k = 0; k < 256; ++k) head[k] = lo[k], tail[k] = lo[k+1];                                        \
	for (k = 0; k < 256;) {                                                                              \
		if (head[k] != tail[k]) {                                                                        \
			int l = (int)((KEY(z[head[k]]) >> shift) & 0xff);                                            \
			if (l != k) {                                                                                \
				TYPE tmp = z[head[k]], swap;                                                             \
				do {                                                                                     \
					swap = tmp; tmp = z[head[l]]; z[head[l]++] = swap;                                   \
					l = (int)((KEY(tmp) >> shift) & 0xff);                                               \
				} while (l != k);                                                                        \
				z[head[k]++] = tmp;                                                                      \
			} else ++head[k];                                                                            \
		} else ++k;                                                                                      \
	}                                                                                                    \
	if (shift) {                                                                                         \
		int next = shift > 8 ? shift - 8 : 0;                                                            \
		for (k = 0; k < 256; ++k) {                                                                      \
			int64_t sz = lo[k+1] - lo[k];                                                                \
			if (sz > ORC_RS_MIN) msd_##NAME(z + lo[k], sz, next);                                        \
			else if (sz > 1) ins_##NAME(z + lo[k], sz);                                                  \
		}                                                                                                \
	}                                                                                                    \
}                                                                                                        \
void orc_radix_sort_##NAME(TYPE *z, int64_t n)                                                           \
{	/* radix_sort_##name, ksort.h:147-151; key is 8 bytes wide => first shift is 56 */                  \
	if (n <= ORC_RS_MIN) ins_##NAME(z, n);                                                               \
	else msd_##NAME(z, n, 56);                                                                           \
}

#define KEY_U64(e) (e)
#define KEY_X(e)   ((e).x)
DEFINE_RADIX(64, uint64_t, KEY_U64)       /* misc.c:139-140 */
DEFINE_RADIX(128x, orc_anchor_t, KEY_X)   /* misc.c:136-137 */

/* ---- tail: chain.c:127-214 ------------------------------------------------------------------------------ */

int32_t orc_bottom(int32_t min_cnt, int32_t min_sc, uint32_t new_i, const orc_seed_t *rec,
                   uint64_t **u_out, orc_anchor_t **b_out, int64_t *n_b_out)
{
	int64_t m = new_i, i, j;
	int32_t n_end = 0, n_keep = 0, n_v = 0, k;
	int32_t *claimed, *members;
	uint64_t *cand, *u, *u2;
	orc_anchor_t *b, *w, *b2;

	*u_out = 0, *b_out = 0, *n_b_out = 0;
	if (m == 0) return 0;
	claimed = (int32_t*)calloc((size_t)m, sizeof(int32_t));
	members = (int32_t*)malloc((size_t)m * sizeof(int32_t));

	/* chain.c:128-135: an end is a record with bit0 set that nobody names as predecessor */
	for (i = 0; i < m; ++i)
		if (rec[i].p >= 0) claimed[rec[i].p >> 2] = 1;
	for (i = 0; i < m; ++i)
		if ((rec[i].p & 1) && !claimed[i]) ++n_end;
	if (n_end == 0) { free(claimed); free(members); return 0; }

	/* chain.c:145-155: walk from each end back to its peak (bit1 == "f < v": not the peak yet) */
	cand = (uint64_t*)malloc((size_t)n_end * 8);
	for (i = 0, n_end = 0; i < m; ++i) {
		if ((rec[i].p & 1) && !claimed[i]) {
			j = i;
			while (j >= 0 && (rec[j].p & 2)) j = rec[j].p >> 2;
			if (j < 0) j = i;
			cand[n_end++] = (uint64_t)rec[j].f << 32 | (uint64_t)j;
		}
	}
	/* chain.c:156-160: ascending radix sort then reverse => descending (peak score, peak index) */
	orc_radix_sort_64(cand, n_end);
	for (i = 0; i < n_end >> 1; ++i) { uint64_t t = cand[i]; cand[i] = cand[n_end-1-i]; cand[n_end-1-i] = t; }

	/* chain.c:163-179: backtrack with claim marks; rejected chains keep their claims */
	memset(claimed, 0, (size_t)m * sizeof(int32_t));
	u = cand; /* overwritten in place exactly like the reference (k <= i always) */
	for (i = 0; i < n_end; ++i) {
		int32_t start_v = n_v, start_k = n_keep, peak_sc = (int32_t)(cand[i] >> 32);
		j = (int32_t)cand[i];
		do {
			members[n_v++] = (int32_t)j;
			claimed[j] = 1;
			j = rec[j].p >> 2;
		} while (j >= 0 && !claimed[j]);
		if (j < 0) {
			if (n_v - start_v >= min_cnt) u[n_keep++] = (uint64_t)peak_sc << 32 | (uint32_t)(n_v - start_v);
		} else if (peak_sc - rec[j].f >= min_sc) {
			if (n_v - start_v >= min_cnt) u[n_keep++] = (uint64_t)(uint32_t)(peak_sc - rec[j].f) << 32 | (uint32_t)(n_v - start_v);
		}
		if (n_keep == start_k) n_v = start_v;
	}
	free(claimed);

	/* chain.c:187-192: anchors of each kept chain in forward order */
	b = (orc_anchor_t*)malloc((size_t)(n_v > 0 ? n_v : 1) * sizeof(orc_anchor_t));
	for (i = 0, k = 0; i < n_keep; ++i) {
		int32_t k0 = k, ni = (int32_t)u[i];
		for (j = 0; j < ni; ++j) b[k++] = rec[members[k0 + (ni - j - 1)]].seed;
	}
	free(members);

	/* chain.c:199-214: order chains by the x of their first anchor with radix_sort_128x */
	w = (orc_anchor_t*)malloc((size_t)(n_keep > 0 ? n_keep : 1) * sizeof(orc_anchor_t));
	for (i = 0, k = 0; i < n_keep; ++i) {
		w[i].x = b[k].x, w[i].y = (uint64_t)k << 32 | (uint64_t)i;
		k += (int32_t)u[i];
	}
	orc_radix_sort_128x(w, n_keep);
	u2 = (uint64_t*)malloc((size_t)(n_keep > 0 ? n_keep : 1) * 8);
	b2 = (orc_anchor_t*)malloc((size_t)(n_v > 0 ? n_v : 1) * sizeof(orc_anchor_t));
	for (i = 0, k = 0; i < n_keep; ++i) {
		int32_t src = (int32_t)w[i].y, cnt = (int32_t)u[src];
		u2[i] = u[src];
		memcpy(&b2[k], &b[w[i].y >> 32], (size_t)cnt * sizeof(orc_anchor_t));
		k += cnt;
	}
	free(w); free(b); free(cand);
	/* NB: the reference returns b even when every candidate was rejected (n_u == 0 after filtering, a
	 * zero-length kmalloc => NULL with kalloc, chain.c:187).  We report n_u = 0 with NULL arrays. */
	if (n_keep == 0) { free(u2); free(b2); return 0; }
	*u_out = u2, *b_out = b2, *n_b_out = n_v;
	return n_keep;
}

int32_t orc_chain_dp(const orc_params_t *prm, int64_t n, const orc_anchor_t *a,
                     uint64_t **u, orc_anchor_t **b, int64_t *n_b)
{
	int32_t *f, *p, *v, n_u;
	orc_seed_t *rec;
	uint32_t new_i;
	*u = 0, *b = 0, *n_b = 0;
	if (n <= 0) return 0;
	f = (int32_t*)malloc((size_t)n * 4); p = (int32_t*)malloc((size_t)n * 4); v = (int32_t*)malloc((size_t)n * 4);
	rec = (orc_seed_t*)malloc((size_t)n * sizeof(orc_seed_t));
	orc_fill(prm, n, a, f, p, v, 0, 0);
	new_i = orc_compact(prm->min_sc, n, a, f, p, v, rec, 0);
	n_u = orc_bottom(prm->min_cnt, prm->min_sc, new_i, rec, u, b, n_b);
	free(f); free(p); free(v); free(rec);
	return n_u;
}

int64_t orc_chain_batch(const orc_params_t *prm, int64_t n_reads, const int64_t *offs, const orc_anchor_t *a,
                        int32_t *n_u, int64_t *u_off, int64_t *b_off, uint64_t **u_out, orc_anchor_t **b_out)
{
	int64_t r, total_vis = 0, cap_u = 1024, cap_b = 4096, tu = 0, tb = 0;
	uint64_t *U = (uint64_t*)malloc((size_t)cap_u * 8);
	orc_anchor_t *B = (orc_anchor_t*)malloc((size_t)cap_b * sizeof(orc_anchor_t));
	for (r = 0; r < n_reads; ++r) {
		int64_t n = offs[r+1] - offs[r], nb = 0;
		const orc_anchor_t *ar = a + offs[r];
		uint64_t *u = 0; orc_anchor_t *b = 0;
		int32_t nu = 0;
		u_off[r] = tu, b_off[r] = tb;
		if (n > 0) {
			int32_t *f = (int32_t*)malloc((size_t)n * 4), *p = (int32_t*)malloc((size_t)n * 4), *v = (int32_t*)malloc((size_t)n * 4);
			orc_seed_t *rec = (orc_seed_t*)malloc((size_t)n * sizeof(orc_seed_t));
			uint32_t new_i;
			total_vis += orc_fill(prm, n, ar, f, p, v, 0, 0);
			new_i = orc_compact(prm->min_sc, n, ar, f, p, v, rec, 0);
			nu = orc_bottom(prm->min_cnt, prm->min_sc, new_i, rec, &u, &b, &nb);
			free(f); free(p); free(v); free(rec);
		}
		n_u[r] = nu;
		if (tu + nu > cap_u) { while (tu + nu > cap_u) cap_u <<= 1; U = (uint64_t*)realloc(U, (size_t)cap_u * 8); }
		if (tb + nb > cap_b) { while (tb + nb > cap_b) cap_b <<= 1; B = (orc_anchor_t*)realloc(B, (size_t)cap_b * sizeof(orc_anchor_t)); }
		if (nu) memcpy(U + tu, u, (size_t)nu * 8);
		if (nb) memcpy(B + tb, b, (size_t)nb * sizeof(orc_anchor_t));
		tu += nu, tb += nb;
		free(u); free(b);
	}
	u_off[n_reads] = tu, b_off[n_reads] = tb;
	*u_out = U, *b_out = B;
	return total_vis;
}

int64_t orc_count_visited(const orc_params_t *prm, int64_t n_reads, const int64_t *offs, const orc_anchor_t *a)
{
	int64_t r, total = 0, cap = 0;
	int32_t *f = 0, *p = 0, *v = 0;
	for (r = 0; r < n_reads; ++r) {
		int64_t n = offs[r+1] - offs[r];
		if (n <= 0) continue;
		if (n > cap) {
			cap = n;
			f = (int32_t*)realloc(f, (size_t)cap * 4); p = (int32_t*)realloc(p, (size_t)cap * 4); v = (int32_t*)realloc(v, (size_t)cap * 4);
		}
		total += orc_fill(prm, n, a + offs[r], f, p, v, 0, 0);
	}
	free(f); free(p); free(v);
	return total;
}

void orc_free(void *p) { free(p); }
