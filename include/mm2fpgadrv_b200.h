/*
 * mm2fpgadrv_b200.h -- the fork's device-driver boundary (SURVEY.md section 8(b) "B2", row f2) served by B200s.
 *
 * stormalex/minimap2_chaindp links a closed libfpgadrv.a; these are the 11 symbols of it that the fork's objects leave
 * undefined (fpga.h:37-62).  libmm2chain_b200.so exports all of them with the same names, signatures and conventions, so the
 * UNMODIFIED fork binary -- every one of its objects, chain.o included -- links against this library in place of
 * libfpgadrv.a and its request packets (minimizers, map.c:286-324) are answered by the GPU: index lookup, seed-hit
 * collection, sort, the DP fill and the compaction run on the device (what map.c:484-575 models on the CPU), and the
 * response packets (new_seed[] + mini_pos[], map.c:533-565) come back through fpga_get_retbuf.
 *
 * Behind the boundary (not the FPGA's structure): submitted packets are queued; a worker thread (MM2B200_SHIM_WORKERS per GPU,
 * default 2, each with its own context and copy of the index streams, so that one batch's packets cross PCIe while the
 * other's is computed) takes EVERYTHING that is waiting (bounded by MM2B200_SHIM_BATCH_MV minimizers) as one device batch.
 * The buffers handed out by fpga_get_writebuf and fpga_get_retbuf are PAGE-LOCKED and recycled through size-classed free
 * lists: a request packet is DMA'd to HBM as the fork wrote it and the minimizers are pulled out of the packet layout on the
 * device; the response packets (headers, records, mini_pos, zero padding) are laid out in HBM and DMA'd into their
 * buffers -- no host thread copies a packet.  (MM2B200_SHIM_ZEROCOPY=0, or a batch the device refuses as a whole, goes the
 * careful way instead: staged and cut on the host, halved on refusal.)  While a batch computes, the fork's worker threads
 * keep submitting, so the batch size adapts to the load without a timer.  With several GPUs (MM2B200_DEVICES=0,1,... or all
 * visible devices with MM2B200_DEVICES=all; default: $MM2B200_DEVICE or 0) the index is replicated and the packets are
 * shared out; reads are independent, there is no collective.
 *
 * No CPU fallback inside the library.  A read the device cannot take (more than 2^21 anchors, a batch of one read that does
 * not fit the GPU) is answered with collect_result_t.err_flag = 1, which is the fork's own protocol for "host, redo this read"
 * (map.c:933-944).  Anything else that fails (no CUDA device, a malformed packet, inconsistent index streams) prints the
 * reason and abort()s, like the entry points of mm2chain_b200.h; MM2B200_ON_ERROR=null answers err_flag = 1 instead.
 */
#ifndef MM2FPGADRV_B200_H
#define MM2FPGADRV_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __FPGA_H__                         /* fpga.h:11-35 -- skip if the reference header is already included */
#define BLOCK        0
#define NOBLOCK      1
#define TYPE_CD      1
#define TYPE_CS      3
#define TYPE_INDEX_B 4
#define TYPE_INDEX_H 5
#define TYPE_INDEX_V 6
#define TYPE_INDEX_P 7
typedef enum { BUF_TYPE_SW = 0, BUF_TYPE_CD = 1, BUF_TYPE_CS = 3 } BUF_TYPE;
typedef enum { RET_TYPE_SW = 0, RET_TYPE_CD = 1, RET_TYPE_CS = 3 } RET_TYPE;

/* fpga.h:37 (main.c:512).  Creates the contexts, the index objects and the worker threads.  0 on success. */
int fpga_init(int flag);
/* fpga.h:39 (main.c:614).  Joins the workers and releases everything; pending packets are finished first. */
void fpga_finalize();

/* fpga.h:58 (main.c:243,581): bw, is_cdna, max_skip, min_sc, opt->flag, opt->mid_occ -- the DP / seeding parameters that are
 * not in the packets.  Takes effect for packets submitted afterwards. */
void fpga_set_params(int bw, int is_cdna, int max_skip, int min_sc, int flag, int max_occ);
/* fpga.h:62 (load_index_buf, main.c:201-204): one chunk of one of the four index streams (TYPE_INDEX_B/H/V/P, written by
 * index.c:603-720).  Waits for packets still in flight, then starts (or continues) a new index; the streams are committed
 * to HBM when the next packet arrives. */
void fpga_load_index(void *addr, int size, int type);

/* fpga.h:44-45 (map.c:439,475; fpga_chaindp.c:105): a host buffer of at least `size` bytes for one request packet, or NULL
 * when MM2B200_SHIM_MAX_BYTES (default 1 GiB) of requests are already waiting -- the caller sleeps and retries
 * (map.c:439-441). */
void *fpga_get_writebuf(unsigned long size, BUF_TYPE type);
void *fpga_get_writebuf_thread(unsigned long size, BUF_TYPE type, int tid);
/* fpga.h:46 (map.c:444,480): hands the filled buffer (a request packet, chaindp_sndhdr_t + tasks) to the device.  The
 * buffer belongs to the library from here on.  0 on success. */
int fpga_writebuf_submit(void *addr, unsigned int size, unsigned int type);

/* fpga.h:41 (fpga_chaindp.c:241): blocks for the next response packet; *len = its size.  After fpga_exit_block(), once
 * nothing is queued or in flight, returns NULL with *len = 0 (the receiver thread's exit signal, fpga_chaindp.c:242-245). */
void *fpga_get_retbuf(int *len, RET_TYPE type);
/* fpga.h:42 (fpga_chaindp.c:262): gives a response buffer back. */
int fpga_release_retbuf(void *addr);

/* fpga.h:52,54 (main.c:608,613): release / re-arm the blocking fpga_get_retbuf. */
void fpga_exit_block();
void fpga_set_block();
#endif

/* Counters of the shim since fpga_init: request packets, reads, device batches, reads answered with err_flag = 1, kernels
 * launched.  Any pointer may be NULL.  Returns 0, or -1 when fpga_init has not run. */
int mm2b200_shim_stats(int64_t *n_packets, int64_t *n_reads, int64_t *n_batches, int64_t *n_err_reads, int64_t *n_launches);

#ifdef __cplusplus
}
#endif
#endif
