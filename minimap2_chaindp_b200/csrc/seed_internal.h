// seed_internal.h -- shared between the seeding kernels (seed_kernels.cu) and their host side (seed_api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "mm2seed_b200.h"

namespace mm2b200 {

// The four index streams of the fork (index.c:603-720) as they lie in HBM.  Nothing is re-encoded: the probe reads the
// same bytes the fork's device would.
struct DevIndex {
	const uint4 *B;        // [2^b] {B[0] lo, B[0] hi, B[1] lo, B[1] hi}
	const uint8_t *H;      // 64 B per group of 8 slots
	const uint64_t *V;     // [n_slots]
	const uint64_t *P;     // [n_pos]
	int32_t b;
	int64_t n_slots, n_pos;
	// Presence filter over the live keys, built once at commit (not part of the fork's streams): two bits per key in a
	// power-of-two bit array of >= 8 bits per hash slot.  A minimizer with a clear bit is absent for sure and never touches
	// B/H; in a noisy read most minimizers are (sequencing errors), and the array is small enough to live in L2.
	const uint32_t *bloom;
	uint64_t bloom_mask;       // number of bits - 1; 0 = no filter
};

__host__ __device__ __forceinline__ uint64_t bloom_bit(uint64_t minier, int which, uint64_t mask)
{
	uint64_t z = minier + (which ? 0x9E3779B97F4A7C15ULL : 0xD1B54A32D192ED03ULL);
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
	return (z ^ (z >> 31)) & mask;
}

enum SeedTotal { ST_ANCHORS = 0, ST_MINI = 1, ST_ERROR = 2, ST_EXACT = 3, ST_MAX_READ = 4, ST_NEXT_READ = 5 /* work counter of k_seed_lookup */, ST_TOTALS = 8 };
enum SeedError { SE_INDEX_RANGE = 1, SE_READ_TOO_LONG = 2 };

#define SEED_IDX_BITS 21                       // anchors of one read are numbered in 21 bits inside the sort key
#define SEED_MAX_READ_ANCHORS (1 << SEED_IDX_BITS)

struct SeedWork {
	// inputs (device copies)
	const ulonglong2 *mv;   // [n_mv] minimizers {x = hash<<8 | span, y = seg<<32 | pos<<1 | strand}
	const int64_t *mv_off;  // [n_reads+1]
	const uint32_t *bid;    // [n_reads]
	const int32_t *qlen;    // [n_reads]
	int32_t n_reads;
	int64_t n_mv;
	int32_t flag, max_occ;
	// per minimizer (written by the lookup)
	uint64_t *src;          // the one position (single-occurrence key) or the index of the first position in P
	int32_t *cnt;           // occurrences to expand (0: absent or repetitive) | 1<<30 single | 1<<31 tandem
	int32_t *aofs;          // offset of the minimizer's first anchor within its read
	int32_t *mread;         // owning read
	uint64_t *mini_sparse;  // mini_pos of read r at mv_off[r] ..
	// per read
	int32_t *n_a, *n_mini, *rep_len;
	int64_t *a_off, *mini_off;     // [n_reads+1]
	uint64_t *mini_out;     // [n_mv] packed
	int64_t *totals;        // [ST_TOTALS]
	// per anchor (allocated once the count is known)
	ulonglong2 *a_un;       // collection order (map.c:200-229)
	ulonglong2 *a;          // sorted (map.c:233)
	uint64_t *key;          // sort keys: order-preserving 43-bit image of x, then the anchor's number within the read
	uint64_t *ky;           // payload copy for the exact replay
	int32_t *stk;           // explicit stack of the exact replay
};

// from chain_api.cu
int api_fail(int code, const char *msg);       // records the thread's last error, returns code
int ctx_device(const mm2b200_ctx *ctx);
int ctx_profile(const mm2b200_ctx *ctx);
int ctx_sm_count(const mm2b200_ctx *ctx);
// per-context state of the stages around the DP (k = 0: post-processing, 1: sketching); freed by mm2b200_destroy
void *ctx_ext(const mm2b200_ctx *ctx, int k);
void ctx_set_ext(mm2b200_ctx *ctx, int k, void *p, void (*free_fn)(void*));
void ctx_add_launches(mm2b200_ctx *ctx, int64_t n);
void sketch_stage_account(mm2b200_ctx *ctx);   // sketch.cu / post.cu: fold the last call's event span into the stage time
void post_stage_account(mm2b200_ctx *ctx);

// launchers; each returns the number of kernels it enqueued
int launch_seed_bloom(const DevIndex &ix, uint32_t *bloom, int32_t *bad, cudaStream_t s);     // fills ix.bloom from the live slots of H; *bad = 1: streams inconsistent
int launch_seed_lookup(const DevIndex &ix, const SeedWork &w, int sm_count, cudaStream_t s);   // + scans over reads, totals, mini_pos packing
int launch_seed_expand(const DevIndex &ix, const SeedWork &w, cudaStream_t s);
int launch_seed_sort(const SeedWork &w, cudaStream_t s);

// ---- the driver shim's device path (seed_api.cu): request packets -> HBM -> response packets, no host copy of a packet
// _begin: the request packets req[p] (req_size[p] bytes; page-locked memory makes each a plain DMA) are laid end to end,
// each padded to 64 bytes; read r has n_mv[r] minimizers at 16 * src16[r] bytes into that concatenation.  Runs
// collect_seed_hits + mm_chain_dp_fpga and returns the per-read counts (page-locked, valid until the next call on the index):
// rec_off[n_reads+1], mini_off[n_reads+1], rep_len[n_reads].  The records and mini_pos stay in HBM for _end.
// _end: resp[p] (resp_size[p] bytes, multiples of 64) laid end to end the same way; read r's records go to byte rec_dst[r]
// and its mini_pos to mini_dst[r] of that concatenation (each area zero-padded to 64 bytes), the n_hdr 64-byte headers in
// hdr to hdr_dst[]; returns when every response packet is in host memory.  Same thread, back to back, one batch at a time.
struct PacketCounts { const int64_t *rec_off, *mini_off; const int32_t *rep_len; int64_t n_anchors; };
int seed_packets_begin(mm2b200_index *ix, const mm2b200_seed_params_t *prm, int n_pkts, const void *const *req, const int64_t *req_size,
                       int64_t n_reads, const int64_t *src16, const int32_t *n_mv, const uint32_t *bid, const int32_t *qlen, PacketCounts *out);
int seed_packets_end(mm2b200_index *ix, int n_pkts, void *const *resp, const int64_t *resp_size, int64_t n_reads,
                     const int64_t *rec_dst, const int64_t *mini_dst, int64_t n_hdr, const void *hdr, const int64_t *hdr_dst);

} // namespace mm2b200
