/*
 * mm2post_b200.h -- chain -> region post-processing on the device, behind the chaining DP.
 *
 * In the reference every read's chains go through mm_gen_regs (hit.c:52-88) and chain_post (map.c:238-247: mm_set_parent
 * hit.c:109-165, mm_select_sub hit.c:230-247, mm_join_long hit.c:290-345) right after mm_chain_dp_bottom
 * (read_result_handle, map.c:837-866).  Here the same steps run on the GPU on the chains where the DP left them, so a
 * host that does not need the anchors (no base-level alignment: the fork's ava mode) downloads a few mm_reg1_t per read
 * instead of every chain anchor.
 *
 * Plain C ABI like mm2chain_b200.h; same error codes, same mm2b200_last_error().  All pointers are HOST pointers unless a
 * name starts with d_.  Single-segment reads only (n_segs <= 1: the long-read presets); multi-segment post-processing
 * (mm_select_sub_multi, mm_seg_gen) is outside this path and rejected with MM2B200_ERR_ARG.
 */
#ifndef MM2POST_B200_H
#define MM2POST_B200_H

#include "mm2chain_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* mm_reg1_t (minimap.h:100-115), bit for bit: 80 bytes.  p (mm_extra_t*, base-alignment results) is always NULL here. */
typedef struct {
	int32_t id, cnt, rid, score;
	int32_t qs, qe, rs, re;
	int32_t parent, subsc;
	int32_t as;                  /* offset of the region's first anchor in the (squeezed) anchor array of its read */
	int32_t mlen, blen;
	int32_t n_sub;
	int32_t score0;
	uint32_t bits;               /* mapq:8, split:2, rev:1, inv:1, sam_pri:1, proper_frag:1, pe_thru:1, seg_split:1, seg_id:8, split_inv:1, dummy:7 */
	uint32_t hash;
	float div;
	void *p;
} mm2b200_reg_t;

#define MM2B200_REG_REV      (1u << 10)   /* bit positions inside mm2b200_reg_t.bits (gcc bit-field layout, x86-64) */
#define MM2B200_REG_INV      (1u << 11)
#define MM2B200_REG_SAM_PRI  (1u << 12)
#define MM2B200_REG_SEG_SPLIT (1u << 15)

/* the mm_mapopt_t fields (minimap.h:124-159) that chain_post and what it calls read */
typedef struct {
	int32_t flag;                /* opt->flag: MM_F_ALL_CHAINS 0x800000 skips everything after mm_gen_regs; MM_F_SPLICE 0x80,
	                                MM_F_SR 0x1000, MM_F_NO_LJOIN 0x400 skip mm_join_long (map.c:240-245) */
	float mask_level;            /* mm_set_parent (hit.c:147) */
	float pri_ratio;             /* mm_select_sub (hit.c:232,238); 0 keeps every hit (ava presets, options.c:87,91) */
	int32_t best_n;              /* mm_select_sub (hit.c:238) */
	int32_t min_diff;            /* mi->k * 2 (map.c:242) */
	int32_t max_join_long, max_join_short, min_join_flank_sc;   /* mm_join_long (hit.c:317-319) */
	int32_t min_cnt;             /* mm_filter_regs after a join (hit.c:255) */
} mm2b200_post_params_t;

/* Results of one batch; arrays are owned by the context (pinned host memory) and valid until its next batch call.
 * Read r has regs[reg_off[r] .. reg_off[r+1]) in the reference's order (hit.c:71-73: by score ^ hash, descending;
 * compacted by mm_select_sub / mm_join_long).  With want_anchors a[a_off[r] .. a_off[r+1]) is the read's chain-anchor array
 * as the reference leaves its a[] (squeezed to the front by mm_squeeze_a when mm_join_long ran, LONG_JOIN marks set,
 * hit.c:265-283,326; what lies behind the last referenced anchor is unspecified); mm2b200_reg_t.as indexes into it. */
typedef struct {
	int64_t n_reads;
	const int64_t *reg_off;          /* [n_reads+1] */
	const mm2b200_reg_t *regs;
	const int64_t *a_off;            /* [n_reads+1]; NULL unless want_anchors */
	const mm128_t *a;
	int64_t n_chains;                /* chains mm_chain_dp produced over the batch (before the regions were selected) */
} mm2b200_regs_t;

/* mm_chain_dp + mm_gen_regs + chain_post over n_reads reads: read r owns a[offs[r] .. offs[r+1]) (sorted by x), hash[r] is
 * the read's name hash (map.c:347-349) and qlen[r] its length (qlen_sum).  Inputs are not consumed.  The batch runs as a
 * pipeline of chunks of about 8 Mi anchors (MM2B200_REGS_CHUNK): the anchors of a chunk cross PCIe while the chunk before
 * is chained and post-processed, and only the regions come back (page-locked a[] makes the upload a plain DMA). */
int mm2b200_chain_regs_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, const mm2b200_post_params_t *post,
                             int64_t n_reads, const int64_t *offs, const mm128_t *a, const uint32_t *hash, const int32_t *qlen,
                             int want_anchors, mm2b200_regs_t *res);

/* Device-resident form.  Post-processes the chains of the LAST mm2b200_chain_device call on ctx (pass its result struct; same
 * stream): d_hash[n_reads] and d_qlen[n_reads] are device pointers; n_chains_cap is an upper bound of the batch's chains
 * (anchors / max(min_cnt, 1) always is one) that sizes the workspace without a device round trip.  The result pointers are
 * valid until the next call on the context; d_totals[0] = number of regions.  The anchors in chains->d_b are updated in
 * place like the reference updates a[] (mm_squeeze_a hit.c:265-283, the LONG_JOIN marks hit.c:326); mm2b200_reg_t.as
 * indexes into the read's slice d_b + d_b_off[r]. */
typedef struct {
	const int64_t *d_reg_off;        /* [n_reads+1] */
	const mm2b200_reg_t *d_regs;
	const int32_t *d_n_regs;         /* [n_reads] */
	const int64_t *d_totals;         /* [8]: [0] regions */
} mm2b200_dev_regs_t;
int mm2b200_post_device(mm2b200_ctx_t *ctx, const mm2b200_post_params_t *post, int64_t n_reads, const mm2b200_dev_result_t *chains,
                        int64_t n_chains_cap, const uint32_t *d_hash, const int32_t *d_qlen, void *stream, mm2b200_dev_regs_t *out);

/* device time of the post-processing stage with cfg.profile = 1 (accumulated since mm2b200_stage_reset) */
int mm2b200_post_stage_ms(mm2b200_ctx_t *ctx, double *ms, int64_t *launches);

#ifdef __cplusplus
}
#endif
#endif
