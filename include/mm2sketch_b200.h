/*
 * mm2sketch_b200.h -- minimizer sketching on the device, in front of the seeding stage.
 *
 * The reference sketches every read on the host (collect_minimizers, map.c:99-111 -> mm_sketch, sketch.c:77-143) and ships
 * the minimizers to its device.  Here the read's bases go to the GPU instead and mm_sketch runs there: the same (w,k)
 * minimizers, in the same order, including the reference's handling of ambiguous bases, symmetric k-mers, equal k-mers in
 * one window and homopolymer compression.  Low-complexity masking (mm_dust_minier, map.c:61-86; off in every preset:
 * sdust_thres = 0, options.c:25) is not part of this path.
 *
 * Plain C ABI like mm2chain_b200.h; same error codes, same mm2b200_last_error().  No CPU fallback.
 */
#ifndef MM2SKETCH_B200_H
#define MM2SKETCH_B200_H

#include "mm2chain_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* argument list of mm_sketch (sketch.c:77) that is constant over a batch */
typedef struct {
	int32_t w;        /* a minimizer for every w consecutive k-mers, 0 < w < 256 (sketch.c:84) */
	int32_t k;        /* k-mer length, 0 < k <= 28 */
	int32_t is_hpc;   /* homopolymer-compressed k-mers (mi->flag & MM_I_HPC, map.c:104) */
} mm2b200_sketch_params_t;

/* Results; arrays are owned by the context (pinned host memory), valid until its next sketch call.
 * Read r has mv[mv_off[r] .. mv_off[r+1]) exactly as mm_sketch appends them: x = hash<<8 | span, y = rid<<32 | lastpos<<1 | strand. */
typedef struct {
	int64_t n_reads;
	const int64_t *mv_off;     /* [n_reads+1] */
	const mm128_t *mv;
} mm2b200_sketch_result_t;

/* mm_sketch over n_reads sequences: read r owns seq[seq_off[r] .. seq_off[r+1]) (bytes as in the FASTA/FASTQ record: any case,
 * anything that is not A/C/G/T/U counts as ambiguous, sketch.c:8-25); rid[r] is mm_sketch's rid argument (NULL = 0 for every
 * read, what collect_minimizers passes for single-segment reads).  Empty reads are allowed and yield nothing.  Inputs are
 * host pointers and are not consumed.  Scratch: 16 bytes of HBM per base of the batch (MM2B200_ERR_NOMEM beyond what the
 * device has: send fewer reads per call). */
int mm2b200_sketch_batch(mm2b200_ctx_t *ctx, const mm2b200_sketch_params_t *prm, int64_t n_reads, const int64_t *seq_off,
                         const char *seq, const uint32_t *rid, mm2b200_sketch_result_t *res);

/* Device-resident form: d_seq / d_seq_off / d_rid (may be NULL) are device pointers, n_bases = seq_off[n_reads] with
 * seq_off[0] = 0.  The work is enqueued on `stream`; the call waits once for the minimizer count (it sizes the packed output),
 * the packing itself is left running on the stream.  d_mv_off[n_reads+1] and d_mv are valid until the next sketch call on the
 * context; d_totals[0] = *n_mv = number of minimizers. */
typedef struct {
	const int64_t *d_mv_off;
	const mm128_t *d_mv;
	const int64_t *d_totals;   /* [8] */
} mm2b200_dev_sketch_t;
int mm2b200_sketch_device(mm2b200_ctx_t *ctx, const mm2b200_sketch_params_t *prm, int64_t n_reads, int64_t n_bases, const int64_t *d_seq_off,
                          const char *d_seq, const uint32_t *d_rid, void *stream, mm2b200_dev_sketch_t *out, int64_t *n_mv);

/* Allocation-free form of mm2b200_sketch_device.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The call allocates nothing, on the host or on the device, and needs no context: the kernels' scratch (16 B per
 * base of minimizer slots, per-read counts) lies in a static device array inside the library, sized for
 * MM2B200_NOALLOC_MAX_READS reads and MM2B200_NOALLOC_MAX_BASES bases per call (16 MiB, held by every GPU that loads the
 * library until the process exits); larger batches are rejected with MM2B200_ERR_ARG.  The device that holds the buffers
 * must be current.  Inputs as for mm2b200_sketch_device (seq_off[0] = 0, seq_off[n_reads] = n_bases).  Outputs are the
 * caller's device buffers: d_mv[mv_cap] (the packed minimizers, same content and order as mm2b200_sketch_device),
 * d_mv_off[n_reads+1], d_totals[8] ([0] = number of minimizers, [1] = 1 when they exceed mv_cap: nothing is written to d_mv
 * then; n_bases is always enough).  Everything is enqueued on `stream`.  With n_mv == NULL the call returns without waiting;
 * otherwise it waits for the stream, stores the count and reports an overflow as MM2B200_ERR_ARG.
 * All calls share the one scratch array: calls that overlap on a GPU (two streams, two threads) race -- serialise them. */
#define MM2B200_NOALLOC_MAX_READS 4096
#define MM2B200_NOALLOC_MAX_BASES (1 << 20)
int mm2b200_sketch_device_noalloc(const mm2b200_sketch_params_t *prm, int64_t n_reads, int64_t n_bases, const int64_t *d_seq_off,
                                  const char *d_seq, const uint32_t *d_rid, void *stream, mm128_t *d_mv, int64_t mv_cap,
                                  int64_t *d_mv_off, int64_t *d_totals, int64_t *n_mv);

/* device time of the sketching stage with cfg.profile = 1 (accumulated) and its kernel launches */
int mm2b200_sketch_stage_ms(mm2b200_ctx_t *ctx, double *ms, int64_t *launches);

#ifdef __cplusplus
}
#endif
#endif
