/*
 * ref_post_harness.c -- thin driver around more UNMODIFIED reference objects (TEST INFRASTRUCTURE ONLY).
 *
 * oracle/Makefile compiles /root/reference/{hit.c,sketch.c,options.c,misc.c,kalloc.c} where they lie and links them with
 * this file into oracle/_ref/libmm2ref_post.so; the reference's own headers are included from $(REF) at compile time.  No
 * reference source is copied into this repository.  The harness makes three things callable on caller-owned arrays:
 *   ref_chain_post  the post-processing sequence of read_result_handle for one read (map.c:862-866): mm_gen_regs, then
 *                   chain_post's calls (map.c:238-247; the function itself is static in map.c, so its five lines of calls are
 *                   replayed here on the reference's own functions)
 *   ref_sketch      mm_sketch (sketch.c:77)
 *   ref_preset      mm_idxopt_init + mm_mapopt_init + mm_set_opt (options.c) -- the presets come from the reference itself
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "minimap.h"
#include "mmpriv.h"
#include "kvec.h"

int32_t mm_idx_cal_max_occ(const mm_idx_t *mi, float f) { (void)mi; (void)f; abort(); return 0; }   /* options.c:mm_mapopt_update, never called here */

typedef struct {          /* same layout as mm2b200_post_params_t / orc_post_params_t */
	int32_t flag; float mask_level, pri_ratio; int32_t best_n, min_diff, max_join_long, max_join_short, min_join_flank_sc, min_cnt;
} ref_post_params_t;

int ref_sizeof_reg(void) { return (int)sizeof(mm_reg1_t); }

/* regs_out must hold n_u records; a[] (the chains' anchors, n_a of them) is modified in place exactly as the reference does
 * (mm_squeeze_a, MM_SEED_LONG_JOIN).  Returns the number of regions. */
int ref_chain_post(const ref_post_params_t *pp, uint32_t hash, int qlen, int n_u, const uint64_t *u, mm128_t *a, mm_reg1_t *regs_out)
{
	mm_mapopt_t opt;
	mm_reg1_t *regs;
	uint64_t *uc;
	int n_regs = n_u;
	if (n_u <= 0) return 0;
	mm_mapopt_init(&opt);
	opt.flag = pp->flag; opt.mask_level = pp->mask_level; opt.pri_ratio = pp->pri_ratio; opt.best_n = pp->best_n;
	opt.max_join_long = pp->max_join_long; opt.max_join_short = pp->max_join_short; opt.min_join_flank_sc = pp->min_join_flank_sc;
	opt.min_cnt = pp->min_cnt;
	uc = (uint64_t*)malloc((size_t)n_u * 8);
	memcpy(uc, u, (size_t)n_u * 8);
	regs = mm_gen_regs(0, hash, qlen, n_u, uc, a);                                 /* map.c:862 */
	if (!(opt.flag & MM_F_ALL_CHAINS)) {                                           /* chain_post, map.c:240-246, n_segs <= 1 */
		mm_set_parent(0, opt.mask_level, n_regs, regs, opt.a * 2 + opt.b);
		mm_select_sub(0, opt.pri_ratio, pp->min_diff, opt.best_n, &n_regs, regs);
		if (!(opt.flag & (MM_F_SPLICE|MM_F_SR|MM_F_NO_LJOIN)))
			mm_join_long(0, &opt, qlen, &n_regs, regs, a);
	}
	memcpy(regs_out, regs, (size_t)n_regs * sizeof(mm_reg1_t));
	free(regs); free(uc);
	return n_regs;
}

/* out must hold at least len + 1 minimizers (mm_sketch emits at most one per base); returns their number */
int64_t ref_sketch(const char *str, int len, int w, int k, uint32_t rid, int is_hpc, mm128_t *out)
{
	mm128_v v = {0, 0, 0};
	int64_t n;
	mm_sketch(0, str, len, w, k, rid, is_hpc, &v);
	n = (int64_t)v.n;
	if (n) memcpy(out, v.a, (size_t)n * sizeof(mm128_t));
	free(v.a);
	return n;
}

typedef struct {
	int32_t k, w, is_hpc, bucket_bits;
	int32_t flag, bw, max_gap, max_gap_ref, max_frag_len, max_chain_skip, min_cnt, min_chain_score;
	float mask_level, pri_ratio; int32_t best_n, max_join_long, max_join_short, min_join_flank_sc;
	float mid_occ_frac; int32_t mid_occ, max_occ;
} ref_preset_t;

/* 0 on success, -1 for an unknown preset name ("" = the defaults of mm_*opt_init) */
int ref_preset(const char *name, ref_preset_t *o)
{
	mm_idxopt_t io;
	mm_mapopt_t mo;
	mm_idxopt_init(&io);
	mm_mapopt_init(&mo);
	if (name && *name && mm_set_opt(name, &io, &mo) < 0) return -1;
	o->k = io.k; o->w = io.w; o->is_hpc = !!(io.flag & MM_I_HPC); o->bucket_bits = io.bucket_bits;
	o->flag = mo.flag; o->bw = mo.bw; o->max_gap = mo.max_gap; o->max_gap_ref = mo.max_gap_ref; o->max_frag_len = mo.max_frag_len;
	o->max_chain_skip = mo.max_chain_skip; o->min_cnt = mo.min_cnt; o->min_chain_score = mo.min_chain_score;
	o->mask_level = mo.mask_level; o->pri_ratio = mo.pri_ratio; o->best_n = mo.best_n;
	o->max_join_long = mo.max_join_long; o->max_join_short = mo.max_join_short; o->min_join_flank_sc = mo.min_join_flank_sc;
	o->mid_occ_frac = mo.mid_occ_frac; o->mid_occ = mo.mid_occ; o->max_occ = mo.max_occ;
	return 0;
}
