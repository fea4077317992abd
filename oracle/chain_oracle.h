/*
 * chain_oracle.h -- CPU restatement of minimap2's anchor-chaining DP (TEST INFRASTRUCTURE ONLY).
 *
 * This is the parity oracle for the B200 path.  It restates, function by function, what
 * /root/reference/chain.c computes, and adds instrumentation the reference does not expose
 * (f/p/v arrays, visited-pair counts).  It is pinned against the compiled reference
 * (oracle/_ref/libmm2ref_chain.so, built by oracle/Makefile from the sources where they lie)
 * and against tests/golden/ -- see tests/test_oracle_vs_ref.py.
 *
 * Nothing under minimap2_chaindp_b200/ may include, link or call this file.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 */
#ifndef CHAIN_ORACLE_H
#define CHAIN_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* anchor: same 16-byte layout as mm128_t (minimap.h:48) */
typedef struct { uint64_t x, y; } orc_anchor_t;

/* compacted record: same 24-byte layout as struct new_seed (minimap.h:51-55) */
typedef struct { orc_anchor_t seed; int32_t p; int32_t f; } orc_seed_t;

/* argument list of mm_chain_dp (chain.c:23) plus the optional max_iter clamp (0 = unbounded = reference) */
typedef struct {
	int32_t max_dist_x, max_dist_y, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs, max_iter;
} orc_params_t;

/* chain.c:46-47 */
float orc_avg_qspan(int64_t n, const orc_anchor_t *a);

/* chain.c:52-90: fill f/p/v (caller-allocated, n each).  visited (may be NULL) gets the number of
 * inner-loop bodies executed for each i; window (may be NULL) gets i - st.  Returns total visited pairs. */
int64_t orc_fill(const orc_params_t *prm, int64_t n, const orc_anchor_t *a,
                 int32_t *f, int32_t *p, int32_t *v, int32_t *visited, int32_t *window);

/* chain.c:93-122: compaction into records; rec has room for n entries; newid (may be NULL, n entries)
 * receives the old->new index map (-1 = dropped).  Returns new_i. */
uint32_t orc_compact(int32_t min_sc, int64_t n, const orc_anchor_t *a,
                     const int32_t *f, const int32_t *p, const int32_t *v,
                     orc_seed_t *rec, int32_t *newid);

/* chain.c:127-214 (== mm_chain_dp_bottom, chain.c:343-430): ends, peaks, ordering, backtrack, gather,
 * x-order.  Outputs are malloc()ed: *u (n_u entries), *b (*n_b anchors).  Returns n_u (0 => *u=*b=NULL). */
int32_t orc_bottom(int32_t min_cnt, int32_t min_sc, uint32_t new_i, const orc_seed_t *rec,
                   uint64_t **u, orc_anchor_t **b, int64_t *n_b);

/* whole of mm_chain_dp (chain.c:23-216) without consuming a[].  Returns n_u. */
int32_t orc_chain_dp(const orc_params_t *prm, int64_t n, const orc_anchor_t *a,
                     uint64_t **u, orc_anchor_t **b, int64_t *n_b);

/* Batch driver used by tests and the CPU-port baseline: read r owns a[offs[r] .. offs[r+1]).
 * Outputs are packed: n_u[r], u (sum n_u), b (sum n_b), with u_off/b_off (n_reads+1 each) caller-allocated.
 * u_out / b_out are malloc()ed by the callee.  Returns total visited pairs (metric numerator). */
int64_t orc_chain_batch(const orc_params_t *prm, int64_t n_reads, const int64_t *offs, const orc_anchor_t *a,
                        int32_t *n_u, int64_t *u_off, int64_t *b_off, uint64_t **u_out, orc_anchor_t **b_out);

/* visited-pair count only (metric numerator), summed over reads */
int64_t orc_count_visited(const orc_params_t *prm, int64_t n_reads, const int64_t *offs, const orc_anchor_t *a);

/* ksort.h:101-151 as instantiated at misc.c:136-140 */
void orc_radix_sort_64(uint64_t *beg, int64_t n);
void orc_radix_sort_128x(orc_anchor_t *beg, int64_t n);

void orc_free(void *p);

#ifdef __cplusplus
}
#endif
#endif
