/*
 * mm2map_b200.h -- the rows of SURVEY.md section 8(f) composed: bases in, regions out, everything between in HBM.
 *
 *   mm_sketch          sketch.c:77-143   (f4, mm2sketch_b200.h)
 *   collect_seed_hits  map.c:187-236     (f1, mm2seed_b200.h) over the index streams of fpga_load_index
 *   mm_chain_dp        chain.c:23        (the hot path, mm2chain_b200.h)
 *   mm_gen_regs + chain_post   hit.c:52-88, map.c:238-247   (f3, mm2post_b200.h)
 *
 * This is mm_map_frag (map.c:326-482 + read_result_handle map.c:816-866) for single-segment reads up to the point where the
 * reference goes on to error estimation, base alignment and mapping quality (mm_est_err, align_regs, mm_set_mapq: host side,
 * outside this path).  One read costs its bases on the way in (1 byte per base) and a few 80-byte mm_reg1_t on the way out;
 * minimizers, anchors and chains never cross PCIe.
 */
#ifndef MM2MAP_B200_H
#define MM2MAP_B200_H

#include "mm2seed_b200.h"
#include "mm2sketch_b200.h"
#include "mm2post_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* arrays are owned by the index object (pinned host memory), valid until its next batch call */
typedef struct {
	int64_t n_reads;
	const int64_t *reg_off;          /* [n_reads+1] */
	const mm2b200_reg_t *regs;       /* mm_reg1_t of read r: regs[reg_off[r] .. reg_off[r+1]) */
	const int32_t *rep_len;          /* [n_reads] (map.c:126-131,142; mm_set_mapq needs it) */
	int64_t n_minimizers, n_anchors, n_chains;
} mm2b200_map_result_t;

/* Read r owns seq[seq_off[r] .. seq_off[r+1]); bid[r] is collect_task_t.bid (rank of the read's name among the target names,
 * bit 31 = the name IS a target name; map.c:262-284,352), hash[r] the read-name hash (map.c:347-349).  seed->chain carries the
 * DP parameters (max_dist_x / max_dist_y = the chaining gaps of map.c:358-366).  Inputs are host pointers, not consumed.
 * The call runs as a pipeline: the bases cross PCIe in parts (page-locked seq[] makes each a plain DMA), the sketching of a
 * part follows its upload on a stream of its own, and the batch is cut into two chunks of whole reads so that the first is
 * seeded, chained and post-processed while the second is still being uploaded and sketched; minimizers, anchors and chains
 * never leave HBM.  Scratch: 16 B of HBM per base of a group of MM2B200_MAP_CHUNK (2 G) bases; larger batches run group by
 * group. */
int mm2b200_map_regs_batch(mm2b200_index_t *idx, const mm2b200_sketch_params_t *sketch, const mm2b200_seed_params_t *seed,
                           const mm2b200_post_params_t *post, int64_t n_reads, const int64_t *seq_off, const char *seq,
                           const uint32_t *bid, const uint32_t *hash, mm2b200_map_result_t *res);

#ifdef __cplusplus
}
#endif
#endif
