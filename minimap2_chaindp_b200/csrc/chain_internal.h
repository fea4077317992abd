// chain_internal.h -- shared between the kernels (chain_kernels.cu) and the host side (chain_api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "mm2chain_b200.h"

namespace mm2b200 {

// One independent anchor run: consecutive x-gaps all <= max_dist_x, so the DP fill of the run
// depends on nothing outside it except the read's avg_qspan.
struct Unit {
	int32_t start;   // index of the first anchor in the batch-wide arrays
	int32_t len;     // >= 2 (singletons are finished by the prep kernel)
	int32_t read;    // owning read (for avg_qspan)
	int32_t pad;
};

enum Counter { CNT_UNITS = 0, CNT_QUEUE = 1, CNT_MULTISEG = 2, CNT_LONG_UNITS = 3, CNT_QUEUE_LONG = 4, CNT_N = 8 };

// Device workspace of one sub-batch ("slot").  All per-anchor arrays are indexed by the batch-wide anchor index;
// per-read scratch of the tail lives at the read's own offset offs[r] (every per-read count is <= n_r).
struct Workspace {
	// inputs
	const ulonglong2 *a;      // anchors (mm128_t) [n_anchors]
	const int64_t *offs;      // [n_reads+1]
	int32_t n_reads, n_anchors;
	// prep
	float *avg;               // [n_reads] avg_qspan (chain.c:47)
	Unit *units;              // [units_cap]: short runs from the front, long runs from the back
	int32_t units_cap;
	int32_t *counters;        // [2 * CNT_N]: the second set belongs to the second half of a split device batch
	int32_t half;             // 0: a whole batch; 1 / 2: this view is the first / second half of a split device batch (mm2b200_chain_device)
	unsigned long long *visited; // [2] reference-visited (i,j) pairs (chain.c:59-88 loop bodies); 32-lane chunks evaluated
	// fill (chain.c:52-90)
	int4 *wrec;               // [n_anchors] working record {low32(x), qpos, f, p}; p is a batch-wide index (or -1)
	int32_t *v;               // [n_anchors] peak score along the best chain ending here (chain.c:90)
	int32_t *f, *p;           // [n_anchors] debug export only (fill_batch with f/p/v requested)
	int32_t *t;               // [n_anchors] marks of the fill that land beyond the current 32-lane chunk (tag = i)
	// compaction (chain.c:93-122)
	int32_t *newid;           // [n_anchors] old -> new index within the read, -1 = dropped
	new_seed *rec;            // [n_anchors] records of read r at rec + offs[r]
	int32_t *recp;            // [n_anchors] copy of rec[].p for the pointer walks
	uint32_t *new_i;          // [n_reads]
	// tail (chain.c:127-214)
	uint8_t *succ, *claimed;  // [n_anchors]
	int32_t *members;         // [n_anchors] v[] of chain.c:168
	uint64_t *cand;           // [n_anchors] u[] before filtering (chain.c:153); reused as sort keys of the x-order
	uint64_t *chain_u;        // [n_anchors] u[] after filtering, score order (chain.c:174-176)
	uint64_t *wy;             // [n_anchors] payload of w[] (chain.c:201)
	int32_t *voff;            // [n_anchors] start of chain k in members[]
	int32_t *order;           // [n_anchors] chain index at final position i
	int32_t *ooff;            // [n_anchors] anchor offset of final position i within the read's output
	int32_t *stk;             // [n_anchors] explicit stack of the radix-sort emulation
	int32_t *n_u;             // [n_reads]
	int32_t *n_v;             // [n_reads]
	// packed outputs
	int64_t *u_off, *b_off;   // [n_reads+1]
	int64_t *totals;          // [8]: sum n_u, sum n_v, n_units, sum new_i, visited pairs
	int64_t *rec_off;         // [n_reads+1] (fill_batch only)
	long long *scan_part;     // [3 * (n_reads/1024 + 2)] per-block sums of the read scans
	uint64_t *u_out;          // [n_anchors]
	ulonglong2 *b_out;        // [n_anchors]
	new_seed *rec_out;        // aliases b_out/u_out region? no: own buffer [n_anchors] (fill_batch only, lazily allocated)
};

struct LaunchCfg {
	int sm_count;
	int fill_blocks_per_sm;      // generic kernel
	int count_visited;           // instrumented fast kernel (metric numerator) instead of the plain one
	int fill_ring;               // ring depth of the fill kernel (0 = chosen per sub-batch from the anchors per read)
};

// launchers (chain_kernels.cu); every launcher returns the number of kernels it enqueued
int launch_prep(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s);
int launch_fill(const Workspace &w, const mm2b200_params_t &prm, const LaunchCfg &cfg, bool generic, cudaStream_t s);
int launch_compact(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s);
int launch_split_records(const Workspace &w, cudaStream_t s);   // bottom_batch: rec[] uploaded by the host
int launch_tail(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s);
int launch_tail_reads(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s);   // per-read part only
int launch_tail_scan(const Workspace &w, cudaStream_t s);                                   // offsets + totals over all reads
int launch_gather(const Workspace &w, cudaStream_t s);
int launch_pack_records(const Workspace &w, cudaStream_t s);
int launch_export_fp(const Workspace &w, cudaStream_t s);       // debug: wrec -> f[], p[] (per-read indices)
// wire format of the upload: [packed u64 x n][block descriptors u64 x blocks][dict u32 x entries]; restores w.a
#define WIRE_BLOCK_SHIFT 8
#define WIRE_BLOCK (1 << WIRE_BLOCK_SHIFT)
int launch_unpack_wire(const Workspace &w, const void *d_wire, size_t tab_pos, size_t dict_pos, cudaStream_t s);
void query_launch_cfg(LaunchCfg *cfg);
long long launch_int_peak(int mode, int grid, int iters, unsigned *d_out, cudaStream_t s);

} // namespace mm2b200
