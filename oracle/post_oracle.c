/*
 * post_oracle.c -- CPU restatement of the reference's chain -> region post-processing.
 * TEST INFRASTRUCTURE ONLY: nothing in the product imports, links or executes this file.
 *
 * Restates, for single-segment reads and without base-level alignment (mm_reg1_t.p == NULL, as on this path):
 *   mm_gen_regs        hit.c:52-88   (with mm_reg_set_coor hit.c:23-38, mm_cal_fuzzy_len hit.c:8-21, hash64 hit.c:40-50)
 *   mm_set_parent      hit.c:109-165
 *   mm_select_sub      hit.c:230-247 (with mm_sync_regs hit.c:207-228, mm_set_sam_pri hit.c:195-205)
 *   mm_join_long       hit.c:290-345 (with mm_squeeze_a hit.c:265-283, mm_filter_regs hit.c:249-263)
 *   chain_post         map.c:238-247 (the order of the calls)
 * Parity: PINNED -- tests/test_post_oracle.py compares every field of every region (and the rewritten anchors) with the
 * compiled reference (oracle/_ref/libmm2ref_post.so) on the golden fixture chains, workload samples and adversarial inputs.
 * The radix sorts are the oracle's own restatements of ksort.h:101-151 (chain_oracle.c), including the unstable path.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "chain_oracle.h"

typedef struct {                     /* mm_reg1_t, minimap.h:100-115 */
	int32_t id, cnt, rid, score, qs, qe, rs, re, parent, subsc, as, mlen, blen, n_sub, score0;
	uint32_t bits, hash;
	float div;
	void *p;
} orc_reg_t;

typedef struct {                     /* same layout as mm2b200_post_params_t */
	int32_t flag; float mask_level, pri_ratio; int32_t best_n, min_diff, max_join_long, max_join_short, min_join_flank_sc, min_cnt;
} orc_post_params_t;

#define PARENT_UNSET   (-1)          /* mmpriv.h:8 */
#define PARENT_TMP_PRI (-2)          /* mmpriv.h:9 */
#define SEED_LONG_JOIN (1ULL<<40)    /* mmpriv.h:16 */
#define F_SPLICE 0x080
#define F_NO_LJOIN 0x400
#define F_SR 0x1000
#define F_ALL_CHAINS 0x800000
#define BIT_REV (1u<<10)
#define BIT_INV (1u<<11)
#define BIT_SAM_PRI (1u<<12)
#define BIT_SEG_SPLIT (1u<<15)

int orc_sizeof_reg(void) { return (int)sizeof(orc_reg_t); }

static uint64_t hash64(uint64_t key)                     /* hit.c:40-50 */
{
	key = (~key + (key << 21));
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8));
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4));
	key = key ^ key >> 28;
	key = (key + (key << 31));
	return key;
}

static void set_coor(orc_reg_t *r, int32_t qlen, const orc_anchor_t *a)      /* hit.c:23-38 + hit.c:8-21 */
{
	int32_t k = r->as, q_span = (int32_t)(a[k].y>>32&0xff), i;
	uint32_t rev = (uint32_t)(a[k].x>>63);
	r->bits = (r->bits & ~BIT_REV) | (rev ? BIT_REV : 0);
	r->rid = (int32_t)(a[k].x<<1>>33);
	r->rs = (int32_t)a[k].x + 1 > q_span? (int32_t)a[k].x + 1 - q_span : 0;
	r->re = (int32_t)a[k + r->cnt - 1].x + 1;
	if (!rev) {
		r->qs = (int32_t)a[k].y + 1 - q_span;
		r->qe = (int32_t)a[k + r->cnt - 1].y + 1;
	} else {
		r->qs = qlen - ((int32_t)a[k + r->cnt - 1].y + 1);
		r->qe = qlen - ((int32_t)a[k].y + 1 - q_span);
	}
	r->mlen = r->blen = 0;
	if (r->cnt <= 0) return;
	r->mlen = r->blen = (int32_t)(a[r->as].y>>32&0xff);
	for (i = r->as + 1; i < r->as + r->cnt; ++i) {
		int span = (int)(a[i].y>>32&0xff);
		int tl = (int32_t)a[i].x - (int32_t)a[i-1].x;
		int ql = (int32_t)a[i].y - (int32_t)a[i-1].y;
		r->blen += tl > ql? tl : ql;
		r->mlen += tl > span && ql > span? span : tl < ql? tl : ql;
	}
}

static void gen_regs(uint32_t hash, int qlen, int n_u, const uint64_t *u, const orc_anchor_t *a, orc_reg_t *r)   /* hit.c:52-88 */
{
	orc_anchor_t *z = (orc_anchor_t*)malloc((size_t)n_u * 16), tmp;
	int i, k;
	for (i = k = 0; i < n_u; ++i) {
		uint32_t h = (uint32_t)hash64((hash64(a[k].x) + hash64(a[k].y)) ^ hash);
		z[i].x = u[i] ^ h;
		z[i].y = (uint64_t)k << 32 | (uint64_t)(int64_t)(int32_t)u[i];
		k += (int32_t)u[i];
	}
	orc_radix_sort_128x(z, n_u);
	for (i = 0; i < n_u>>1; ++i) tmp = z[i], z[i] = z[n_u-1-i], z[n_u-1-i] = tmp;
	memset(r, 0, (size_t)n_u * sizeof(orc_reg_t));
	for (i = 0; i < n_u; ++i) {
		orc_reg_t *ri = &r[i];
		ri->id = i;
		ri->parent = PARENT_UNSET;
		ri->score = ri->score0 = (int32_t)(z[i].x >> 32);
		ri->hash = (uint32_t)z[i].x;
		ri->cnt = (int32_t)z[i].y;
		ri->as = (int32_t)(z[i].y >> 32);
		ri->div = -1.0f;
		set_coor(ri, qlen, a);
	}
	free(z);
}

static void set_parent(float mask_level, int n, orc_reg_t *r)               /* hit.c:109-165 with p == NULL */
{
	int i, j, k, *w;
	uint64_t *cov;
	if (n <= 0) return;
	for (i = 0; i < n; ++i) r[i].id = i;
	cov = (uint64_t*)malloc((size_t)n * 8);
	w = (int*)malloc((size_t)n * sizeof(int));
	w[0] = 0, r[0].parent = 0;
	for (i = 1, k = 1; i < n; ++i) {
		orc_reg_t *ri = &r[i];
		int si = ri->qs, ei = ri->qe, n_cov = 0, uncov_len = 0;
		for (j = 0; j < k; ++j) {
			orc_reg_t *rp = &r[w[j]];
			int sj = rp->qs, ej = rp->qe;
			if (ej <= si || sj >= ei) continue;
			if (sj < si) sj = si;
			if (ej > ei) ej = ei;
			cov[n_cov++] = (uint64_t)sj<<32 | (uint64_t)(int64_t)ej;
		}
		if (n_cov > 0) {
			int x = si;
			orc_radix_sort_64(cov, n_cov);
			for (j = 0; j < n_cov; ++j) {
				if (cov[j]>>32 > (uint64_t)(int64_t)x) uncov_len = (int)((uint64_t)(int64_t)uncov_len + ((cov[j]>>32) - (uint64_t)(int64_t)x));
				x = (int32_t)cov[j] > x? (int32_t)cov[j] : x;
			}
			if (ei > x) uncov_len += ei - x;
			for (j = 0; j < k; ++j) {
				orc_reg_t *rp = &r[w[j]];
				int sj = rp->qs, ej = rp->qe, min, max, ol;
				if (ej <= si || sj >= ei) continue;
				min = ej - sj < ei - si? ej - sj : ei - si;
				max = ej - sj > ei - si? ej - sj : ei - si;
				ol = si < sj? (ei < sj? 0 : ei < ej? ei - sj : ej - sj) : (ej < si? 0 : ej < ei? ej - si : ei - si);
				if ((float)ol / min - (float)uncov_len / max > mask_level) {
					ri->parent = rp->parent;
					rp->subsc = rp->subsc > ri->score? rp->subsc : ri->score;
					if (ri->cnt >= rp->cnt) ++rp->n_sub;
					break;
				}
			}
		} else j = k;
		if (j == k) w[k++] = i, ri->parent = i, ri->n_sub = 0;
	}
	free(cov); free(w);
}

static void set_sam_pri(int n, orc_reg_t *r)                                 /* hit.c:195-205 */
{
	int i, n_pri = 0;
	for (i = 0; i < n; ++i) {
		r[i].bits &= ~BIT_SAM_PRI;
		if (r[i].id == r[i].parent) {
			++n_pri;
			if (n_pri == 1) r[i].bits |= BIT_SAM_PRI;
		}
	}
}

static void sync_regs(int n_regs, orc_reg_t *regs)                           /* hit.c:207-228 */
{
	int *tmp, i, max_id = -1, n_tmp;
	if (n_regs <= 0) return;
	for (i = 0; i < n_regs; ++i) max_id = max_id > regs[i].id? max_id : regs[i].id;
	n_tmp = max_id + 1;
	tmp = (int*)malloc((size_t)(n_tmp > 0 ? n_tmp : 1) * sizeof(int));
	for (i = 0; i < n_tmp; ++i) tmp[i] = -1;
	for (i = 0; i < n_regs; ++i)
		if (regs[i].id >= 0) tmp[regs[i].id] = i;
	for (i = 0; i < n_regs; ++i) {
		orc_reg_t *r = &regs[i];
		r->id = i;
		if (r->parent == PARENT_TMP_PRI) r->parent = i;
		else if (r->parent >= 0 && tmp[r->parent] >= 0) r->parent = tmp[r->parent];
		else r->parent = PARENT_UNSET;
	}
	free(tmp);
	set_sam_pri(n_regs, regs);
}

static void select_sub(float pri_ratio, int min_diff, int best_n, int *n_, orc_reg_t *r)   /* hit.c:230-247 */
{
	if (pri_ratio > 0.0f && *n_ > 0) {
		int i, k, n = *n_, n_2nd = 0;
		for (i = k = 0; i < n; ++i) {
			int p = r[i].parent;
			if (p == i || (r[i].bits & BIT_INV)) {
				r[k++] = r[i];
			} else if ((r[i].score >= r[p].score * pri_ratio || r[i].score + min_diff >= r[p].score) && n_2nd < best_n) {
				if (!(r[i].qs == r[p].qs && r[i].qe == r[p].qe && r[i].rid == r[p].rid && r[i].rs == r[p].rs && r[i].re == r[p].re))
					r[k++] = r[i], ++n_2nd;
			}
		}
		if (k != n) sync_regs(k, r);
		*n_ = k;
	}
}

static void filter_regs(int min_cnt, int *n_regs, orc_reg_t *regs)            /* hit.c:249-263 with p == NULL */
{
	int i, k;
	for (i = k = 0; i < *n_regs; ++i) {
		orc_reg_t *r = &regs[i];
		int flt = 0;
		if (!(r->bits & BIT_INV) && !(r->bits & BIT_SEG_SPLIT) && r->cnt < min_cnt) flt = 1;
		if (!flt) {
			if (k < i) regs[k++] = regs[i];
			else ++k;
		}
	}
	*n_regs = k;
}

static void squeeze_a(int n_regs, orc_reg_t *regs, orc_anchor_t *a)            /* hit.c:265-283 */
{
	int i, as = 0;
	uint64_t *aux = (uint64_t*)malloc((size_t)n_regs * 8);
	for (i = 0; i < n_regs; ++i) aux[i] = (uint64_t)regs[i].as << 32 | (uint64_t)i;
	orc_radix_sort_64(aux, n_regs);
	for (i = 0; i < n_regs; ++i) {
		orc_reg_t *r = &regs[(int32_t)aux[i]];
		if (r->as != as) {
			memmove(&a[as], &a[r->as], (size_t)r->cnt * 16);
			r->as = as;
		}
		as += r->cnt;
	}
	free(aux);
}

static void join_long(const orc_post_params_t *opt, int qlen, int *n_regs_, orc_reg_t *regs, orc_anchor_t *a)   /* hit.c:290-345 */
{
	int i, n_aux, n_regs = *n_regs_, n_drop = 0;
	uint64_t *aux;
	if (n_regs < 2) return;
	squeeze_a(n_regs, regs, a);
	aux = (uint64_t*)malloc((size_t)n_regs * 8);
	for (i = n_aux = 0; i < n_regs; ++i)
		if (regs[i].parent == i || regs[i].parent < 0)
			aux[n_aux++] = (uint64_t)regs[i].as << 32 | (uint64_t)i;
	orc_radix_sort_64(aux, n_aux);
	for (i = n_aux - 1; i >= 1; --i) {
		orc_reg_t *r0 = &regs[(int32_t)aux[i-1]], *r1 = &regs[(int32_t)aux[i]];
		orc_anchor_t *a0e, *a1s;
		int max_gap, min_gap, sc_thres;
		if (r0->as + r0->cnt != r1->as) continue;
		if (r0->rid != r1->rid || (r0->bits & BIT_REV) != (r1->bits & BIT_REV)) continue;
		a0e = &a[r0->as + r0->cnt - 1];
		a1s = &a[r1->as];
		if (a1s->x <= a0e->x || (int32_t)a1s->y <= (int32_t)a0e->y) continue;
		max_gap = min_gap = (int32_t)a1s->y - (int32_t)a0e->y;
		max_gap = (int)((uint64_t)(int64_t)max_gap > a1s->x - a0e->x? (uint64_t)(int64_t)max_gap : a1s->x - a0e->x);
		min_gap = (int)((uint64_t)(int64_t)min_gap < a1s->x - a0e->x? (uint64_t)(int64_t)min_gap : a1s->x - a0e->x);
		if (max_gap > opt->max_join_long || min_gap > opt->max_join_short) continue;
		sc_thres = (int)((float)opt->min_join_flank_sc / opt->max_join_long * max_gap + .499);
		if (r0->score < sc_thres || r1->score < sc_thres) continue;
		if (r0->re - r0->rs < max_gap>>1 || r0->qe - r0->qs < max_gap>>1) continue;
		if (r1->re - r1->rs < max_gap>>1 || r1->qe - r1->qs < max_gap>>1) continue;
		a[r1->as].y |= SEED_LONG_JOIN;
		r0->cnt += r1->cnt, r0->score += r1->score;
		set_coor(r0, qlen, a);
		r1->cnt = 0;
		r1->parent = r0->id;
		++n_drop;
	}
	free(aux);
	if (n_drop > 0) {
		for (i = 0; i < n_regs; ++i) {
			orc_reg_t *r = &regs[i];
			if (r->parent >= 0 && r->id != r->parent) {
				if (regs[r->parent].parent >= 0 && regs[r->parent].parent != r->parent)
					r->parent = regs[r->parent].parent;
			}
		}
		filter_regs(opt->min_cnt, n_regs_, regs);
		sync_regs(*n_regs_, regs);
	}
}

/* the sequence of read_result_handle (map.c:862-866) for one read; a[] is rewritten in place like the reference does */
int orc_chain_post(const orc_post_params_t *pp, uint32_t hash, int qlen, int n_u, const uint64_t *u, orc_anchor_t *a, orc_reg_t *regs)
{
	int n_regs = n_u;
	if (n_u <= 0) return 0;
	gen_regs(hash, qlen, n_u, u, a, regs);
	if (!(pp->flag & F_ALL_CHAINS)) {                                          /* chain_post, map.c:240-246 */
		set_parent(pp->mask_level, n_regs, regs);
		select_sub(pp->pri_ratio, pp->min_diff, pp->best_n, &n_regs, regs);
		if (!(pp->flag & (F_SPLICE|F_SR|F_NO_LJOIN))) join_long(pp, qlen, &n_regs, regs, a);
	}
	return n_regs;
}
