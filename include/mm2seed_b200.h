/*
 * mm2seed_b200.h -- seed-hit collection on the device, in front of the chaining DP.
 *
 * In the fork, a request packet carries a read's MINIMIZERS (not its anchors) to the device; the device owns a copy of
 * the index and does collect_seed_hits + the DP (map.c:484-575 is the fork's software model of that device).  This
 * header is the same split for the B200: the index streams are uploaded once, a batch call takes minimizers and returns
 * chains, and the anchors never cross PCIe.
 *
 * Plain C ABI like mm2chain_b200.h; same error codes, same mm2b200_last_error().  All pointers are HOST pointers.
 */
#ifndef MM2SEED_B200_H
#define MM2SEED_B200_H

#include "mm2chain_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- the index on the device ------------------------------------------------------------------------------- */

/* stream types of fpga_load_index (fpga.h:20-23 TYPE_INDEX_B/H/V/P; written by index.c:603-720) */
#define MM2B200_INDEX_B 4   /* 16 B per bucket: offsets into H and P, size of the bucket's hash table */
#define MM2B200_INDEX_H 5   /* hash tables: 64 B per 8 slots = flags word, 8 x 48-bit keys, padding */
#define MM2B200_INDEX_V 6   /* hash values, 8 B per slot */
#define MM2B200_INDEX_P 7   /* position lists */

typedef struct mm2b200_index mm2b200_index_t;

/* An empty index bound to ctx (its device, its chaining workspace).  One batch at a time per index object. */
int  mm2b200_index_create(mm2b200_ctx_t *ctx, mm2b200_index_t **idx);
void mm2b200_index_destroy(mm2b200_index_t *idx);

/* Replaces fpga_load_index(addr, size, type) (fpga.h:62; called from load_index_buf, main.c:201-204): appends one
 * chunk of a stream.  Chunks of one type arrive in order; types may interleave. */
int mm2b200_index_load(mm2b200_index_t *idx, const void *addr, int64_t size, int type);

/* Ends the upload: checks the stream sizes against each other (16 B x 2^b buckets; 8 values per 64-byte group),
 * copies the streams to HBM and releases the host copies.  Must precede the first batch; a later mm2b200_index_load
 * starts a new index. */
int mm2b200_index_commit(mm2b200_index_t *idx);

/* Forgets the stream chunks loaded since the last commit and the committed index: the next mm2b200_index_load starts from
 * nothing.  (The driver shim calls it when a new index begins on an object that never served a packet of the old one.) */
int mm2b200_index_reset(mm2b200_index_t *idx);

/* log2(number of buckets) (mi->b) and the bytes resident in HBM, after commit */
int mm2b200_index_info(const mm2b200_index_t *idx, int32_t *b, int64_t *bytes);

/* ---- seeding + chaining over a batch of reads ----------------------------------------------------------------- */

/* what fpga_set_params(bw, is_cdna, max_skip, min_sc, flag, max_occ) (fpga.h:58; main.c:243) gives the device, plus the
 * per-packet fields that are constant over a run (collect_task_t.gap_ref / gap_qry, map.c:358-366,394-395) */
typedef struct {
	mm2b200_params_t chain;   /* max_dist_x = gap_ref, max_dist_y = gap_qry, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs */
	int32_t flag;             /* opt->flag: MM_F_NO_DIAG 0x1, MM_F_NO_DUAL 0x2, MM_F_FOR_ONLY 0x100000, MM_F_REV_ONLY 0x200000 are read (map.c:146-185) */
	int32_t max_occ;          /* opt->mid_occ: minimizers with >= max_occ occurrences count as repetitive (map.c:126) */
} mm2b200_seed_params_t;

/* Results; arrays are owned by the index object (pinned host memory), valid until its next batch call.
 *   chains      as mm2b200_chain_batch would return for the anchors collect_seed_hits yields (map.c:187-236)
 *   rep_len[r]  length of the query covered by repetitive minimizers (map.c:126-131,142)
 *   mini_pos[mini_off[r] .. mini_off[r+1])  q_span<<32 | q_pos>>1 of the non-repetitive minimizers (map.c:139)
 *   a[a_off[r] .. a_off[r+1])  the sorted anchors themselves -- only when want_anchors != 0 (tests, debugging) */
typedef struct {
	mm2b200_result_t chains;
	int64_t n_anchors;
	const int32_t  *rep_len;    /* [n_reads] */
	const int64_t  *mini_off;   /* [n_reads+1] */
	const uint64_t *mini_pos;
	const int64_t  *a_off;      /* [n_reads+1] (always filled) */
	const mm128_t  *a;          /* NULL unless want_anchors */
	int64_t n_exact_sorts;      /* reads whose anchors tie in x and exceed 64: their (unstable) reference order was replayed */
} mm2b200_seed_result_t;

/* Read r owns minimizers mv[mv_off[r] .. mv_off[r+1]) exactly as mm_sketch emits them and the request packet carries
 * them (map.c:286-324; sketch.c: x = hash<<8 | span, y = seg<<32 | pos<<1 | strand); bid[r] and qlen[r] are collect_task_t.bid
 * and .qlensum (fpga_chaindp.h:46-58).  Inputs are not consumed. */
int mm2b200_seed_chain_batch(mm2b200_index_t *idx, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off,
                             const mm128_t *mv, const uint32_t *bid, const int32_t *qlen, int want_anchors,
                             mm2b200_seed_result_t *res);

/* Same work, but the answer is what the fork's device sends back (collect_result_t + new_seed[] + mini_pos[], map.c:533-565):
 * per read the compacted DP records of mm_chain_dp_fpga instead of finished chains. */
typedef struct {
	int64_t n_anchors;
	const int64_t *rec_off;          /* [n_reads+1] */
	const struct new_seed *rec;
	const int32_t  *rep_len;
	const int64_t  *mini_off;
	const uint64_t *mini_pos;
} mm2b200_seed_records_t;
int mm2b200_seed_fill_batch(mm2b200_index_t *idx, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off,
                            const mm128_t *mv, const uint32_t *bid, const int32_t *qlen, mm2b200_seed_records_t *res);

/* Packet level: the request packet of map.c:286-324 in, the response packet of map.c:484-575 out (fpga_writebuf_submit ->
 * fpga_get_retbuf in the fork).  resp must hold resp_cap bytes; *resp_size receives the bytes written.  Reads of one
 * packet share gap_ref / gap_qry in the fork; packets that mix them are split into groups internally.  n_segs is read from
 * gap_qry like the reference's model does (map.c:517). */
int mm2b200_seed_work(mm2b200_index_t *idx, const mm2b200_seed_params_t *prm, const void *req, int64_t req_size,
                      void *resp, int64_t resp_cap, int64_t *resp_size);

/* kernels launched by the seeding stage of this index object so far, and device time of its stages with cfg.profile = 1 */
int64_t mm2b200_seed_launch_count(const mm2b200_index_t *idx);
int mm2b200_seed_stage_ms(mm2b200_index_t *idx, double *lookup_ms, double *expand_ms, double *sort_ms);

#ifdef __cplusplus
}
#endif
#endif
