/*
 * mm2chain_b200.h -- C ABI of libmm2chain_b200.so: minimap2's anchor-chaining DP on NVIDIA B200 (sm_100a).
 *
 * The library is a drop-in for ONE path of the reference (stormalex/minimap2_chaindp, minimap2 2.10-r761):
 * chain.c.  Citations are file:line into the reference tree.
 *
 *   Section 1  the reference's own entry points, same names / signatures / ownership (mmpriv.h:68-70,
 *              chain.c:23,218,329) -- map.c links against this library instead of chain.o.
 *   Section 2  context + pinned "kalloc-style" arenas (the GPU analogue of km_init / the fork's
 *              fpga_get_writebuf, fpga.h:44-46).
 *   Section 3  the batched host API the reference-side batcher calls (many reads per call; replaces the
 *              per-packet fpga_writebuf_submit / fpga_get_retbuf round trip, map.c:423-444, fpga_chaindp.c:228-270).
 *   Section 4  device-resident entry points (inputs already in HBM; used for kernel-only timing) and
 *              parity/debug hooks that expose f[]/p[]/v[].
 *
 * All functions return MM2B200_OK (0) or a negative error code; mm2b200_last_error() gives the message for the
 * calling thread.  Nothing here aborts.  There is no CPU fallback: without a usable CUDA device every compute
 * entry point fails with MM2B200_ERR_CUDA.
 */
#ifndef MM2CHAIN_B200_H
#define MM2CHAIN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- types shared with the reference (bit-for-bit) -------------------------------------------------------- */

#ifndef MINIMAP2_H                       /* minimap.h:48 -- skip if the reference header is already included */
typedef struct { uint64_t x, y; } mm128_t;   /* x = strand<<63 | rid<<32 | rpos ; y = seg<<48 | flags | q_span<<32 | qpos */
struct new_seed {                        /* minimap.h:51-55: compacted DP record */
	mm128_t seed;
	int32_t p;                           /* (new index of predecessor, or -1)<<2 | (f<v)<<1 | (v>=min_sc) */
	int32_t f;                           /* score of the best chain ending here */
};
#endif

/* argument list of mm_chain_dp (chain.c:23).  max_iter is an extension (upstream >= 2.12): 0 = unbounded,
 * which is the behaviour of this reference; >0 clamps the predecessor window to the last max_iter anchors. */
typedef struct {
	int32_t max_dist_x, max_dist_y, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs, max_iter;
} mm2b200_params_t;

#define MM2B200_OK             0
#define MM2B200_ERR_CUDA      -1   /* CUDA runtime error / no device */
#define MM2B200_ERR_ARG       -2   /* bad argument (negative sizes, unsorted offsets, n >= 2^29 per read, ...) */
#define MM2B200_ERR_NOMEM     -3   /* host or device allocation failed */
#define MM2B200_ERR_STATE     -4   /* context used concurrently / after destroy */

/* ---- Section 1: reference entry points --------------------------------------------------------------------- */

/* chain.c:23 (mmpriv.h:68).  Consumes a (kfree(km,a)); returns kmalloc(km)ed b and *_u; NULL + *n_u_=0 + *_u=0
 * when there is no chain.  Thread-safe: concurrent callers are combined into one device batch by the
 * library's batcher (default context on device $MM2B200_DEVICE, default 0).  kmalloc/kfree are taken from the
 * host program when it exports them (kalloc.c:120,72); otherwise km must be NULL and malloc/free are used.
 * Errors (no CUDA device, out of memory, n >= 2^29, ...): the reference has no error channel and "NULL, 0" also means
 * "no chain", so a failing drop-in symbol prints the reason to stderr and abort()s (default).  With the environment
 * variable MM2B200_ON_ERROR=null it returns NULL with *n_u_ = 0 instead and mm2b200_last_error() is non-empty (a is
 * still consumed).  A failing batch is retried call by call, so one bad read cannot fail its neighbours. */
mm128_t *mm_chain_dp(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_cnt, int min_sc, int is_cdna,
                     int n_segs, int64_t n, mm128_t *a, int *n_u_, uint64_t **_u, void *km);

/* chain.c:218 (mmpriv.h:69): the fork's device half.  free()s a; returns a malloc()ed new_seed[*_new_i].
 * Thread-safe and batched like mm_chain_dp (the fork's live call sites are map.c:525,941). */
struct new_seed *mm_chain_dp_fpga(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_sc, int is_cdna,
                                  int n_segs, int64_t n, mm128_t *a, uint32_t *_new_i);

/* chain.c:329 (mmpriv.h:70): the fork's host half (here also on the device).  Does not free fpga_a.
 * Returns malloc()ed b and *_u (the reference ignores km here, chain.c:341-358,400).  Thread-safe and batched like
 * mm_chain_dp (the fork's live call site is map.c:837). */
mm128_t *mm_chain_dp_bottom(int min_cnt, int min_sc, int n_segs, int *n_u_, uint64_t **_u, void *km,
                            struct new_seed *fpga_a, uint32_t new_i);

/* how many device batches the three entry points above were combined into, and how many calls they served */
int mm2b200_entry_stats(int64_t *n_batches, int64_t *n_calls);

/* ---- Section 2: context ------------------------------------------------------------------------------------ */

typedef struct mm2b200_ctx mm2b200_ctx_t;

typedef struct {
	int32_t device;              /* CUDA device ordinal */
	int32_t n_slots;             /* in-flight sub-batches (streams) for H2D / compute / D2H overlap; 0 = default (3) */
	int64_t slot_anchors;        /* anchors per sub-batch; 0 = default (8 Mi).  A longer read grows its slot. */
	int32_t profile;             /* 1: record CUDA events around every kernel (see mm2b200_stage_ms) */
	int32_t count_visited;       /* 1: the fill kernel also counts the (i,j) pairs the reference loop visits
	                                (mm2b200_result_t.n_visited, the throughput metric's numerator); off by default */
} mm2b200_config_t;

int  mm2b200_create(const mm2b200_config_t *cfg, mm2b200_ctx_t **ctx);
void mm2b200_destroy(mm2b200_ctx_t *ctx);
const char *mm2b200_last_error(void);
int  mm2b200_device_count(void);

/* page-locked host memory the batcher can DMA from directly (any other host pointer is staged through the
 * context's own pinned buffers) */
void *mm2b200_pinned_alloc(size_t bytes);
void  mm2b200_pinned_free(void *p);

/* ---- Section 3: batched host API --------------------------------------------------------------------------- */

/* Results of one batch.  All arrays are owned by the context (pinned host memory) and stay valid until the next
 * batch call on the same context or mm2b200_destroy.  Read r produced n_u[r] chains:
 *   u[u_off[r] + k] = score<<32 | n_anchors   (chain.c:174-176), chains ordered by first-anchor x (chain.c:199-212)
 *   b[b_off[r] ...]  = the chains' anchors, each chain in ascending order (chain.c:187-192, 206-213)          */
typedef struct {
	int64_t n_reads;
	const int32_t  *n_u;     /* [n_reads] */
	const int64_t  *u_off;   /* [n_reads+1] */
	const int64_t  *b_off;   /* [n_reads+1] */
	const uint64_t *u;       /* [u_off[n_reads]] */
	const mm128_t  *b;       /* [b_off[n_reads]] */
	int64_t n_units;         /* independent anchor runs (len >= 2) the fill kernel processed */
	int64_t n_sub_batches;
	int64_t n_visited;       /* (i,j) pairs the reference loop visits (bodies of chain.c:59-88); needs cfg.count_visited */
	int64_t n_lanes;         /* predecessor lanes the fill kernel evaluated (32 per chunk); n_visited / n_lanes = lane efficiency */
} mm2b200_result_t;

/* mm_chain_dp over n_reads reads: read r owns a[offs[r] .. offs[r+1]) (each sorted by x, map.c:233).
 * a and offs are host pointers and are NOT consumed. */
int mm2b200_chain_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *offs,
                        const mm128_t *a, mm2b200_result_t *res);

/* mm_chain_dp_fpga over a batch: rec[rec_off[r] .. rec_off[r+1]) are read r's new_seed records (context-owned).
 * If f/p/v are non-NULL (host arrays of offs[n_reads] int32) the raw DP arrays are copied out too, with p[]
 * holding per-read indices exactly like the reference's p[] (chain.c:89). */
typedef struct {
	int64_t n_reads;
	const int64_t *rec_off;          /* [n_reads+1] */
	const struct new_seed *rec;      /* [rec_off[n_reads]] */
} mm2b200_records_t;
int mm2b200_fill_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *offs,
                       const mm128_t *a, mm2b200_records_t *out, int32_t *f, int32_t *p, int32_t *v);

/* mm_chain_dp_bottom over a batch of record arrays (rec_off[n_reads+1], host pointers, not consumed) */
int mm2b200_bottom_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *rec_off,
                         const struct new_seed *rec, mm2b200_result_t *res);

/* ---- Section 3b: read-batch sharding over several GPUs of one box ------------------------------------------- */

/* Splits reads [0, n_reads) into n_parts contiguous ranges of near-equal anchor count (the cost model of SURVEY 8e:
 * work ~ anchors).  cuts[n_parts+1] receives the range boundaries (cuts[0] = 0, cuts[n_parts] = n_reads).
 * Pure host arithmetic: usable without a GPU. */
int mm2b200_partition_reads(int64_t n_reads, const int64_t *offs, int32_t n_parts, int64_t *cuts);

/* mm_chain_dp over a batch sharded across n_ctx contexts (normally one per GPU): one host thread per context,
 * no collective -- reads are independent.  Shard k = reads [cuts[k], cuts[k+1]) (cuts[n_ctx+1], filled by the call);
 * res[k] is that shard's result, owned by ctxs[k], with offsets relative to the shard.  Results stay in submission
 * order: shard 0's reads first. */
int mm2b200_chain_batch_multi(mm2b200_ctx_t *const *ctxs, int32_t n_ctx, const mm2b200_params_t *prm, int64_t n_reads,
                              const int64_t *offs, const mm128_t *a, int64_t *cuts, mm2b200_result_t *res);

/* ---- Section 4: device-resident entry points, instrumentation ---------------------------------------------- */

/* Device pointers valid until the next call on the context.  Sizes live on the device (totals[0] = sum n_u,
 * totals[1] = sum of chain lengths, totals[2] = n_units) so the call never synchronises with the host. */
typedef struct {
	const int32_t  *d_n_u;    /* [n_reads] */
	const int64_t  *d_u_off;  /* [n_reads+1] */
	const int64_t  *d_b_off;  /* [n_reads+1] */
	const uint64_t *d_u;
	const mm128_t  *d_b;
	const int64_t  *d_totals; /* [8]: [0..2] as above, [3] = sum new_i, [4] = visited (i,j) pairs, [5] = 32-lane chunks evaluated */
} mm2b200_dev_result_t;

/* Whole path on anchors already resident in HBM: d_a[n_anchors] and d_offs[n_reads+1] are device pointers.
 * Everything is enqueued on `stream` (a cudaStream_t passed as void*; NULL = the legacy default stream) and the
 * call returns without waiting.  The device result pointers stay valid until the next call on the context; the library
 * records an event behind the work, and a later call on another stream (or one that has to grow the workspace) waits
 * for it before it touches the slot. */
int mm2b200_chain_device(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, int64_t n_anchors,
                         const int64_t *d_offs, const mm128_t *d_a, void *stream, mm2b200_dev_result_t *out);

/* mm_chain_dp_fpga on anchors already resident in HBM: like mm2b200_chain_device, but stops after the compaction
 * (chain.c:93-122).  d_rec[d_rec_off[r] .. d_rec_off[r+1]) are read r's new_seed records, d_totals[3] = their number. */
typedef struct {
	const int64_t *d_rec_off;        /* [n_reads+1] */
	const struct new_seed *d_rec;
	const int64_t *d_totals;         /* [8], as in mm2b200_dev_result_t */
} mm2b200_dev_records_t;
int mm2b200_fill_device(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, int64_t n_anchors,
                        const int64_t *d_offs, const mm128_t *d_a, void *stream, mm2b200_dev_records_t *out);

/* mm2b200_chain_device on >= 4096 reads runs the two halves of the batch on two streams, so that the HBM-bound stages
 * of one half run beside the issue-bound fill of the other (default on; MM2B200_SPLIT=0 at context creation turns it off).
 * With on = 0 every kernel of a pass runs alone on the caller's stream -- the setting under which a kernel's own duration
 * (mm2b200_stage_ms, the roofline of bench.py) is measured. */
int mm2b200_set_split(mm2b200_ctx_t *ctx, int32_t on);

/* kernel launches issued by this context so far (bench.py's gpu_launches) */
int64_t mm2b200_launch_count(const mm2b200_ctx_t *ctx);

/* with cfg.profile = 1: accumulated device time (ms, CUDA events on the launching stream) and launch count of a
 * stage since the last reset.  Stage names: "prep", "fill", "compact", "tail", "gather". */
int  mm2b200_stage_ms(mm2b200_ctx_t *ctx, const char *stage, double *ms, int64_t *launches);
void mm2b200_stage_reset(mm2b200_ctx_t *ctx);

/* Measured integer-issue peak of the device (lane-ops/s): the roofline denominator for the predecessor scan.
 * mode 0 = ALU-pipe only (add/xor), mode 1 = ALU + FMA-pipe integer ops interleaved (the issue-port ceiling). */
int mm2b200_int_peak(mm2b200_ctx_t *ctx, int mode, double *lane_ops_per_s);

/* Host batcher (north_star subsystem 1; the reference's counterpart is the packet builder fpga_chaindp.c:send_task).
 * With wire != 0 (opt-in; the default is the raw 16-byte upload) every sub-batch of anchors is encoded by `threads` host
 * threads (0 = all cores, at most 16 by default) into a lossless wire format of 8 bytes per anchor plus small per-block dictionaries before it crosses
 * PCIe, and is restored to mm128_t on the device.  Per block of 256 consecutive anchors the high half of x (strand|rid)
 * and the high half of y (flags|seg|qspan) are replaced by dictionary indices; a sub-batch holding a block whose qpos
 * values do not fit the 24 bits left (minus the bits of the block's y index) is uploaded raw.  Results are identical
 * either way.  It pays only where the host can stream anchors through its cores faster than PCIe moves them (not on the
 * measured 16-vCPU box, see profiles/).  Environment at context creation: MM2B200_WIRE=1, MM2B200_WIRE_THREADS=n. */
int mm2b200_set_upload(mm2b200_ctx_t *ctx, int32_t wire, int32_t threads);
/* The encoder alone (host only, no CUDA): writes [packed u64 x n][block descriptors u64 x blocks][dict u32 x entries]
 * into buf and reports the layout.  Returns 1 when the anchors do not fit the format.  For tests and tools. */
int mm2b200_wire_encode(const mm128_t *a, int64_t n, void *buf, size_t cap, size_t *bytes, size_t *tab_pos, size_t *dict_pos);
/* bytes copied host -> device by the last batch call on this context, and how many of its sub-batches used the wire format */
int mm2b200_upload_stats(const mm2b200_ctx_t *ctx, int64_t *h2d_bytes, int64_t *wire_sub_batches);

const char *mm2b200_version(void);

#ifdef __cplusplus
}
#endif
#endif
