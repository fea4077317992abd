// seed_api.cu -- host side of the seeding stage (include/mm2seed_b200.h): the index streams in HBM, the batch calls that go
// minimizers -> anchors -> chains without the anchors ever leaving the device, and the packet-level entry that speaks the
// fork's request / response format (map.c:286-324, 484-575).
//
// No CPU fallback: every call needs the device the context was created on.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <time.h>
#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <vector>
#include "seed_internal.h"
#include "packet_format.h"
#include "mm2map_b200.h"

namespace mm2b200 {
// sketch.cu: mm_sketch for a group of chunks that share the slot arrays; reads [part_cut[k], part_cut[k+1]) of a chunk are in HBM
// once part_ev[k] has fired -- kernels of a chunk on the sketching stream, its scan
// and packing later on the compute stream
int sketch_pipe_launch(mm2b200_ctx *ctx, const mm2b200_sketch_params_t *prm, bool first, bool last, int64_t group_reads, int64_t group_bases, int group_parts,
                       int64_t r_off, int64_t b_off, int part_off, int64_t nr, const int64_t *d_seq_off, const char *d_seq, cudaStream_t st,
                       int n_parts, const int64_t *part_cut, const cudaEvent_t *part_ev, cudaEvent_t done_ev);
int sketch_pipe_finish(mm2b200_ctx *ctx, int64_t r_off, int64_t b_off, int64_t nr, const int64_t *d_seq_off, cudaStream_t st, cudaEvent_t done_ev,
                       mm2b200_dev_sketch_t *out, int64_t *n_mv);
}
using namespace mm2b200;

namespace {

int failf(int code, const char *fmt, ...) __attribute__((format(printf, 2, 3)));
int failf(int code, const char *fmt, ...)
{
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	return api_fail(code, buf);
}

#define CU_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
	return failf(MM2B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

inline size_t up256(size_t x) { return (x + 255) / 256 * 256; }

// MM2B200_TRACE=1: host wall-clock marks of the composed path on stderr (ms since the call began)
struct Trace {
	bool on; struct timespec t0;
	Trace() { const char *e = getenv("MM2B200_TRACE"); on = e && atoi(e); if (on) clock_gettime(CLOCK_MONOTONIC, &t0); }
	void mark(const char *what) const
	{
		if (!on) return;
		struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t);
		fprintf(stderr, "[mm2b200 trace] %-28s %8.3f ms\n", what, 1e3 * (double)(t.tv_sec - t0.tv_sec) + 1e-6 * (double)(t.tv_nsec - t0.tv_nsec));
	}
};

struct DevBuf {
	void *p = nullptr; size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap && p) return MM2B200_OK;
		size_t want = bytes + bytes / 4 + 4096;
		if (p) { cudaFree(p); p = nullptr; cap = 0; }
		if (cudaMalloc(&p, want) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaMalloc(%zu bytes) failed in the seeding stage", want); }
		cap = want;
		return MM2B200_OK;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct PinBuf {
	void *p = nullptr; size_t cap = 0;
	int reserve(size_t bytes)
	{
		if (bytes <= cap && p) return MM2B200_OK;
		size_t want = bytes + bytes / 4 + 4096;
		if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
		if (cudaHostAlloc(&p, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu bytes) failed in the seeding stage", want); }
		cap = want;
		return MM2B200_OK;
	}
	// grows like reserve but keeps the first `keep` bytes (the caller has synchronised everything that touches the buffer)
	int reserve_keep(size_t bytes, size_t keep)
	{
		if (bytes <= cap && p) return MM2B200_OK;
		const size_t want = bytes + bytes / 2 + 4096;
		void *q = nullptr;
		if (cudaHostAlloc(&q, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu bytes) failed in the seeding stage", want); }
		if (p && keep) memcpy(q, p, keep);
		if (p) cudaFreeHost(p);
		p = q; cap = want;
		return MM2B200_OK;
	}
	void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
	template <class T> T *as() { return (T*)p; }
};

// carves one device allocation into arrays
struct Carver {
	struct Item { void **pp; size_t bytes; };
	std::vector<Item> items;
	template <class T> void add(T **pp, size_t bytes) { items.push_back(Item{(void**)pp, bytes}); }
	size_t total() const { size_t t = 0; for (auto &it : items) t += up256(it.bytes); return t; }
	void place(void *base) { char *cur = (char*)base; for (auto &it : items) { *it.pp = cur; cur += up256(it.bytes); } }
};

} // namespace

struct mm2b200_index {
	mm2b200_ctx_t *ctx = nullptr;
	int device = 0;
	std::mutex mu;
	// upload
	std::vector<uint8_t> host[4];          // B, H, V, P chunks appended until commit
	bool committed = false;
	DevBuf d_stream[4], d_bloom;
	size_t l2_window = 0; float l2_ratio = 0;  // presence filter kept in L2 while the lookup runs (0: not available / switched off)
	DevIndex dix{};
	int64_t resident_bytes = 0;
	// per-batch state
	cudaStream_t stream = nullptr;
	DevBuf d_in, d_work, d_anchor;
	PinBuf h_stage, h_totals, h_rep, h_mini_off, h_mini, h_a_off, h_a;
	PinBuf h_n_u, h_u_off, h_b_off, h_u, h_b, h_rec_off, h_rec;
	DevBuf d_reads;                        // mm2b200_map_regs_batch: bases, offsets, bid, hash, qlen
	PinBuf h_reg_off, h_regs, h_chunk_off;
	cudaStream_t copy_stream = nullptr;
	std::vector<cudaEvent_t> ev_chunk;
	cudaEvent_t ev_small = nullptr;
	cudaStream_t sk_stream = nullptr;       // mm2b200_map_regs_batch: the sketching kernels follow the upload on their own stream
	std::vector<cudaEvent_t> ev_sk;         // one per chunk: its parts are sketched
	// packet path of the driver shim (seed_packets_begin / _end): raw request packets, gathered minimizers, response packets
	DevBuf d_pkt_in, d_pkt_mv, d_pkt_meta, d_pkt_out;
	PinBuf h_pkt, h_pkt_cnt;
	const int64_t *pk_rec_off = nullptr, *pk_mini_off = nullptr; const void *pk_rec = nullptr; const uint64_t *pk_mini = nullptr;
	int64_t pk_reads = -1, pk_rec_total = 0;
	int64_t launches = 0;
	cudaEvent_t ev[4] = { nullptr, nullptr, nullptr, nullptr };
	double ms[3] = { 0, 0, 0 };
};

extern "C" int mm2b200_index_create(mm2b200_ctx_t *ctx, mm2b200_index_t **out)
{
	if (!ctx || !out) return failf(MM2B200_ERR_ARG, "NULL ctx or index pointer");
	*out = nullptr;
	mm2b200_index *ix = new mm2b200_index();
	ix->ctx = ctx;
	ix->device = ctx_device(ctx);
	cudaError_t e = cudaSetDevice(ix->device);
	if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking);
	for (int k = 0; k < 4 && e == cudaSuccess; ++k) e = cudaEventCreate(&ix->ev[k]);
	if (e != cudaSuccess) {
		int rc = failf(MM2B200_ERR_CUDA, "index_create: %s", cudaGetErrorString(e));
		delete ix;
		return rc;
	}
	*out = ix;
	return MM2B200_OK;
}

extern "C" void mm2b200_index_destroy(mm2b200_index_t *ix)
{
	if (!ix) return;
	cudaSetDevice(ix->device);
	if (ix->stream) cudaStreamSynchronize(ix->stream);
	for (auto &b : ix->d_stream) b.release();
	ix->d_bloom.release();
	ix->d_in.release(); ix->d_work.release(); ix->d_anchor.release(); ix->d_reads.release();
	ix->h_reg_off.release(); ix->h_regs.release(); ix->h_chunk_off.release();
	ix->d_pkt_in.release(); ix->d_pkt_mv.release(); ix->d_pkt_meta.release(); ix->d_pkt_out.release(); ix->h_pkt.release(); ix->h_pkt_cnt.release();
	if (ix->copy_stream) { cudaStreamSynchronize(ix->copy_stream); cudaStreamDestroy(ix->copy_stream); }
	for (auto e : ix->ev_chunk) cudaEventDestroy(e);
	if (ix->ev_small) cudaEventDestroy(ix->ev_small);
	if (ix->sk_stream) { cudaStreamSynchronize(ix->sk_stream); cudaStreamDestroy(ix->sk_stream); }
	for (auto e : ix->ev_sk) cudaEventDestroy(e);
	for (PinBuf *b : { &ix->h_stage, &ix->h_totals, &ix->h_rep, &ix->h_mini_off, &ix->h_mini, &ix->h_a_off, &ix->h_a,
	                   &ix->h_n_u, &ix->h_u_off, &ix->h_b_off, &ix->h_u, &ix->h_b, &ix->h_rec_off, &ix->h_rec }) b->release();
	for (auto &e : ix->ev) if (e) cudaEventDestroy(e);
	if (ix->stream) cudaStreamDestroy(ix->stream);
	delete ix;
}

extern "C" int mm2b200_index_load(mm2b200_index_t *ix, const void *addr, int64_t size, int type)
{
	if (!ix) return failf(MM2B200_ERR_ARG, "NULL index");
	if (type < MM2B200_INDEX_B || type > MM2B200_INDEX_P) return failf(MM2B200_ERR_ARG, "index stream type %d is not one of B/H/V/P (4..7)", type);
	if (size < 0 || (size > 0 && !addr)) return failf(MM2B200_ERR_ARG, "bad index chunk");
	std::lock_guard<std::mutex> lk(ix->mu);
	if (ix->committed) {            // a new index replaces the old one
		ix->committed = false;
		for (auto &h : ix->host) h.clear();
	}
	std::vector<uint8_t> &h = ix->host[type - MM2B200_INDEX_B];
	try { h.insert(h.end(), (const uint8_t*)addr, (const uint8_t*)addr + size); }
	catch (const std::bad_alloc &) { return failf(MM2B200_ERR_NOMEM, "out of host memory while buffering an index stream"); }
	return MM2B200_OK;
}

extern "C" int mm2b200_index_reset(mm2b200_index_t *ix)
{
	if (!ix) return failf(MM2B200_ERR_ARG, "NULL index");
	std::lock_guard<std::mutex> lk(ix->mu);
	ix->committed = false;
	for (auto &h : ix->host) std::vector<uint8_t>().swap(h);
	return MM2B200_OK;
}

extern "C" int mm2b200_index_commit(mm2b200_index_t *ix)
{
	if (!ix) return failf(MM2B200_ERR_ARG, "NULL index");
	std::lock_guard<std::mutex> lk(ix->mu);
	const size_t nB = ix->host[0].size(), nH = ix->host[1].size(), nV = ix->host[2].size(), nP = ix->host[3].size();
	int b = 0;
	while (((size_t)16 << b) < nB && b < 40) ++b;
	if (nB < 16 || ((size_t)16 << b) != nB) return failf(MM2B200_ERR_ARG, "B stream of %zu bytes is not 16 bytes x a power of two of buckets", nB);
	if (nH % 64 || nV % 8 || nP % 8 || nV / 8 != nH / 64 * 8)
		return failf(MM2B200_ERR_ARG, "index streams disagree: H %zu bytes (64 per 8 slots), V %zu bytes (8 per slot), P %zu bytes", nH, nV, nP);
	CU_TRY(cudaSetDevice(ix->device));
	CU_TRY(cudaStreamSynchronize(ix->stream));
	ix->resident_bytes = 0;
	for (int k = 0; k < 4; ++k) {
		const size_t n = ix->host[k].size();
		int rc = ix->d_stream[k].reserve(n ? n : 256);
		if (rc) return rc;
		if (n) CU_TRY(cudaMemcpy(ix->d_stream[k].p, ix->host[k].data(), n, cudaMemcpyHostToDevice));
		ix->resident_bytes += (int64_t)n;
		std::vector<uint8_t>().swap(ix->host[k]);
	}
	ix->dix.B = (const uint4*)ix->d_stream[0].p;
	ix->dix.H = (const uint8_t*)ix->d_stream[1].p;
	ix->dix.V = (const uint64_t*)ix->d_stream[2].p;
	ix->dix.P = (const uint64_t*)ix->d_stream[3].p;
	ix->dix.b = b;
	ix->dix.n_slots = (int64_t)(nV / 8);
	ix->dix.n_pos = (int64_t)(nP / 8);
	ix->dix.bloom = nullptr; ix->dix.bloom_mask = 0; ix->l2_window = 0;
	{	// presence filter over the live keys (seed_internal.h): >= 8 bits per hash slot, at least 2^20, at most 2^33 bits
		const char *eb = getenv("MM2B200_SEED_BLOOM");
		if (!(eb && atoi(eb) == 0) && ix->dix.n_slots > 0) {
			uint64_t bits = (uint64_t)1 << 20;
			while (bits < (uint64_t)ix->dix.n_slots * 8 && bits < ((uint64_t)1 << 33)) bits <<= 1;
			int rc = ix->d_bloom.reserve(bits / 8 + 256);          // + a flag word behind the bits
			if (rc) return rc;
			CU_TRY(cudaMemsetAsync(ix->d_bloom.p, 0, bits / 8 + 256, ix->stream));
			int32_t *d_bad = (int32_t*)((char*)ix->d_bloom.p + bits / 8), bad = 0;
			ix->dix.bloom_mask = bits - 1;
			ix->dix.bloom = (const uint32_t*)ix->d_bloom.p;
			ix->launches += launch_seed_bloom(ix->dix, (uint32_t*)ix->d_bloom.p, d_bad, ix->stream);
			CU_TRY(cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, ix->stream));
			CU_TRY(cudaStreamSynchronize(ix->stream));
			CU_TRY(cudaGetLastError());
			if (bad) { ix->dix.bloom = nullptr; ix->dix.bloom_mask = 0; }      // a bucket points outside H: the lookup will say so
			else {
				ix->resident_bytes += (int64_t)(bits / 8);
				// The filter is probed at random by every minimizer while the lookup streams gigabytes of minimizers and results
				// through L2: marked persisting for that kernel (a carve-out of L2 no streaming line can evict), its probes stop
				// going to HBM.  MM2B200_SEED_L2PIN=0 switches this off.
				const char *ep = getenv("MM2B200_SEED_L2PIN");
				cudaDeviceProp prop;
				if (!(ep && atoi(ep) == 0) && cudaGetDeviceProperties(&prop, ix->device) == cudaSuccess && prop.persistingL2CacheMaxSize > 0 && prop.accessPolicyMaxWindowSize > 0) {
					const size_t bytes = (size_t)(bits / 8);
					const size_t carve = bytes < (size_t)prop.persistingL2CacheMaxSize ? bytes : (size_t)prop.persistingL2CacheMaxSize;
					size_t cur = 0;
					cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
					if (cur >= carve || cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess) {
						ix->l2_window = bytes < (size_t)prop.accessPolicyMaxWindowSize ? bytes : (size_t)prop.accessPolicyMaxWindowSize;
						ix->l2_ratio = ix->l2_window <= carve ? 1.0f : (float)carve / (float)ix->l2_window;
					}
					cudaGetLastError();
				}
			}
		}
	}
	ix->committed = true;
	return MM2B200_OK;
}

extern "C" int mm2b200_index_info(const mm2b200_index_t *ix, int32_t *b, int64_t *bytes)
{
	if (!ix || !ix->committed) return failf(MM2B200_ERR_STATE, "index not committed");
	if (b) *b = ix->dix.b;
	if (bytes) *bytes = ix->resident_bytes;
	return MM2B200_OK;
}

extern "C" int64_t mm2b200_seed_launch_count(const mm2b200_index_t *ix) { return ix ? ix->launches : 0; }

extern "C" int mm2b200_seed_stage_ms(mm2b200_index_t *ix, double *lookup_ms, double *expand_ms, double *sort_ms)
{
	if (!ix) return failf(MM2B200_ERR_ARG, "NULL index");
	if (lookup_ms) *lookup_ms = ix->ms[0];
	if (expand_ms) *expand_ms = ix->ms[1];
	if (sort_ms) *sort_ms = ix->ms[2];
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// minimizers -> sorted anchors in HBM
// ---------------------------------------------------------------------------------------------------------
namespace {

struct SeedOut {
	SeedWork w{};
	int64_t n_anchors = 0, n_mini = 0, n_exact = 0;
};

int check_seed_args(const mm2b200_index *ix, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off, const mm128_t *mv,
                    const uint32_t *bid, const int32_t *qlen)
{
	if (!ix) return failf(MM2B200_ERR_ARG, "NULL index");
	if (!ix->committed) return failf(MM2B200_ERR_STATE, "index not committed (mm2b200_index_commit)");
	if (!prm) return failf(MM2B200_ERR_ARG, "params is NULL");
	if (n_reads < 0 || n_reads >= ((int64_t)1 << 31) - 1) return failf(MM2B200_ERR_ARG, "bad number of reads");
	if (n_reads && (!mv_off || !bid || !qlen)) return failf(MM2B200_ERR_ARG, "NULL per-read array");
	for (int64_t r = 0; r < n_reads; ++r)
		if (mv_off[r + 1] < mv_off[r]) return failf(MM2B200_ERR_ARG, "mv_off is not non-decreasing at read %lld", (long long)r);
	if (n_reads && mv_off[n_reads] > mv_off[0] && !mv) return failf(MM2B200_ERR_ARG, "NULL minimizer array");
	return MM2B200_OK;
}

// uploads the batch, runs lookup -> (sync for the anchor count) -> expand -> sort on ix->stream.  On return the sorted
// anchors and their offsets are in HBM (out.w.a, out.w.a_off) and the stream holds the remaining work.
// dev_n_mv < 0: mv_off / mv / bid / qlen are host arrays (uploaded here).  dev_n_mv >= 0: they are DEVICE arrays holding
// dev_n_mv minimizers with mv_off[0] = 0 (the sketching stage left them in HBM); host_meta = false then also skips the
// download of rep_len / mini_pos / offsets (the caller takes what it needs from out.w).
int seed_stage(mm2b200_index *ix, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off, const mm128_t *mv,
               const uint32_t *bid, const int32_t *qlen, SeedOut &out, int64_t dev_n_mv = -1, bool host_meta = true)
{
	int rc;
	CU_TRY(cudaSetDevice(ix->device));
	cudaStream_t st = ix->stream;
	const bool on_device = dev_n_mv >= 0;
	const int64_t m_base = on_device ? 0 : (n_reads ? mv_off[0] : 0);
	const int64_t n_mv = on_device ? dev_n_mv : (n_reads ? mv_off[n_reads] - m_base : 0);
	const size_t R = (size_t)n_reads, M = (size_t)n_mv;
	SeedWork &w = out.w;
	// inputs
	ulonglong2 *d_mv = nullptr; int64_t *d_mv_off = nullptr; uint32_t *d_bid = nullptr; int32_t *d_qlen = nullptr;
	if (!on_device) {
		Carver c;
		c.add(&d_mv, 16 * (M + 1)); c.add(&d_mv_off, 8 * (R + 1)); c.add(&d_bid, 4 * (R + 1)); c.add(&d_qlen, 4 * (R + 1));
		if ((rc = ix->d_in.reserve(c.total()))) return rc;
		c.place(ix->d_in.p);
	}
	{
		Carver c;
		c.add(&w.src, 8 * (M + 1)); c.add(&w.cnt, 4 * (M + 1)); c.add(&w.aofs, 4 * (M + 1)); c.add(&w.mread, 4 * (M + 1));
		c.add(&w.mini_sparse, 8 * (M + 1)); c.add(&w.mini_out, 8 * (M + 1));
		c.add(&w.n_a, 4 * (R + 1)); c.add(&w.n_mini, 4 * (R + 1)); c.add(&w.rep_len, 4 * (R + 1));
		c.add(&w.a_off, 8 * (R + 1)); c.add(&w.mini_off, 8 * (R + 1)); c.add(&w.totals, 8 * ST_TOTALS);
		if ((rc = ix->d_work.reserve(c.total()))) return rc;
		c.place(ix->d_work.p);
	}
	if (on_device) {
		w.mv = (const ulonglong2*)mv; w.mv_off = mv_off; w.bid = bid; w.qlen = qlen;
	} else {
		// offsets relative to the first minimizer of the batch, staged in pinned memory together with bid / qlen
		if ((rc = ix->h_stage.reserve(8 * (R + 1) + 8 * (R + 1)))) return rc;
		int64_t *h_off = ix->h_stage.as<int64_t>();
		uint32_t *h_bid = (uint32_t*)(h_off + (R + 1));
		int32_t *h_qlen = (int32_t*)(h_bid + (R + 1));
		for (size_t r = 0; r <= R; ++r) h_off[r] = n_reads ? mv_off[r] - m_base : 0;
		if (R) { memcpy(h_bid, bid, 4 * R); memcpy(h_qlen, qlen, 4 * R); }
		CU_TRY(cudaMemcpyAsync(d_mv_off, h_off, 8 * (R + 1), cudaMemcpyHostToDevice, st));
		if (R) {
			CU_TRY(cudaMemcpyAsync(d_bid, h_bid, 4 * R, cudaMemcpyHostToDevice, st));
			CU_TRY(cudaMemcpyAsync(d_qlen, h_qlen, 4 * R, cudaMemcpyHostToDevice, st));
		}
		if (M) CU_TRY(cudaMemcpyAsync(d_mv, mv + m_base, 16 * M, cudaMemcpyHostToDevice, st));
		w.mv = d_mv; w.mv_off = d_mv_off; w.bid = d_bid; w.qlen = d_qlen;
	}
	w.n_reads = (int32_t)n_reads; w.n_mv = n_mv;
	w.flag = prm->flag; w.max_occ = prm->max_occ;
	const bool prof = ctx_profile(ix->ctx) != 0;
	if (prof) CU_TRY(cudaEventRecord(ix->ev[0], st));
	if (ix->l2_window) {
		cudaStreamAttrValue av;
		memset(&av, 0, sizeof(av));
		av.accessPolicyWindow.base_ptr = (void*)ix->dix.bloom; av.accessPolicyWindow.num_bytes = ix->l2_window;
		av.accessPolicyWindow.hitRatio = ix->l2_ratio; av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting; av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
		if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av) != cudaSuccess) { cudaGetLastError(); ix->l2_window = 0; }
	}
	ix->launches += launch_seed_lookup(ix->dix, w, ctx_sm_count(ix->ctx), st);
	if (ix->l2_window) {                           // the kernels behind the lookup get an ordinary L2 again
		cudaStreamAttrValue av;
		memset(&av, 0, sizeof(av));
		cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av);
		cudaGetLastError();
	}
	if (prof) CU_TRY(cudaEventRecord(ix->ev[1], st));
	if ((rc = ix->h_totals.reserve(8 * ST_TOTALS * 2))) return rc;
	int64_t *tot = ix->h_totals.as<int64_t>();
	CU_TRY(cudaMemcpyAsync(tot, w.totals, 8 * ST_TOTALS, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	CU_TRY(cudaGetLastError());
	if (tot[ST_ERROR] & SE_INDEX_RANGE) return failf(MM2B200_ERR_ARG, "index streams are inconsistent: a bucket or position list points outside H/V/P");
	if (tot[ST_ERROR] & SE_READ_TOO_LONG)
		return failf(MM2B200_ERR_ARG, "a read has %lld anchors; the seeding stage numbers a read's anchors in %d bits", (long long)tot[ST_MAX_READ], SEED_IDX_BITS);
	const int64_t A = tot[ST_ANCHORS];
	if (A >= ((int64_t)1 << 31) - 64) return failf(MM2B200_ERR_ARG, "batch yields %lld anchors: beyond the 2^31 index space of one sub-batch, send fewer reads per call", (long long)A);
	out.n_anchors = A; out.n_mini = tot[ST_MINI];
	{
		Carver c;
		const size_t a = (size_t)A + 1;
		c.add(&w.a_un, 16 * a); c.add(&w.a, 16 * a); c.add(&w.key, 8 * a); c.add(&w.ky, 8 * a); c.add(&w.stk, 4 * a + 64);
		if ((rc = ix->d_anchor.reserve(c.total()))) return rc;
		c.place(ix->d_anchor.p);
	}
	if (A) {
		ix->launches += launch_seed_expand(ix->dix, w, st);
		if (prof) CU_TRY(cudaEventRecord(ix->ev[2], st));
		ix->launches += launch_seed_sort(w, st);
		if (prof) CU_TRY(cudaEventRecord(ix->ev[3], st));
	}
	CU_TRY(cudaGetLastError());
	if (!host_meta) return MM2B200_OK;
	// what the host always gets from this stage: rep_len, mini_pos, anchor offsets
	if ((rc = ix->h_rep.reserve(4 * (R + 1))) || (rc = ix->h_mini_off.reserve(8 * (R + 1))) || (rc = ix->h_a_off.reserve(8 * (R + 1))) ||
	    (rc = ix->h_mini.reserve(8 * ((size_t)out.n_mini + 1)))) return rc;
	if (R) CU_TRY(cudaMemcpyAsync(ix->h_rep.p, w.rep_len, 4 * R, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(ix->h_mini_off.p, w.mini_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(ix->h_a_off.p, w.a_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	if (out.n_mini) CU_TRY(cudaMemcpyAsync(ix->h_mini.p, w.mini_out, 8 * (size_t)out.n_mini, cudaMemcpyDeviceToHost, st));
	return MM2B200_OK;
}

// after the stream has been synchronised: stage times and the exact-sort counter
int seed_finish(mm2b200_index *ix, SeedOut &out)
{
	if (ctx_profile(ix->ctx)) {
		float t = 0;
		if (cudaEventElapsedTime(&t, ix->ev[0], ix->ev[1]) == cudaSuccess) ix->ms[0] += t;
		if (out.n_anchors) {
			if (cudaEventElapsedTime(&t, ix->ev[1], ix->ev[2]) == cudaSuccess) ix->ms[1] += t;   // includes the host's wait for the count
			if (cudaEventElapsedTime(&t, ix->ev[2], ix->ev[3]) == cudaSuccess) ix->ms[2] += t;
		}
		cudaGetLastError();
	}
	return MM2B200_OK;
}

} // namespace

// ---------------------------------------------------------------------------------------------------------
// batch calls
// ---------------------------------------------------------------------------------------------------------
extern "C" int mm2b200_seed_chain_batch(mm2b200_index_t *ix, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off,
                                        const mm128_t *mv, const uint32_t *bid, const int32_t *qlen, int want_anchors,
                                        mm2b200_seed_result_t *res)
{
	int rc;
	if (!res) return failf(MM2B200_ERR_ARG, "NULL result");
	if ((rc = check_seed_args(ix, prm, n_reads, mv_off, mv, bid, qlen))) return rc;
	std::lock_guard<std::mutex> lk(ix->mu);
	memset(res, 0, sizeof(*res));
	SeedOut so;
	if ((rc = seed_stage(ix, prm, n_reads, mv_off, mv, bid, qlen, so))) return rc;
	cudaStream_t st = ix->stream;
	const size_t R = (size_t)n_reads, A = (size_t)so.n_anchors;
	if (want_anchors) {
		if ((rc = ix->h_a.reserve(16 * (A + 1)))) return rc;
		if (A) CU_TRY(cudaMemcpyAsync(ix->h_a.p, so.w.a, 16 * A, cudaMemcpyDeviceToHost, st));
	}
	// the chaining DP reads the anchors where the sort left them
	mm2b200_dev_result_t dev;
	if ((rc = mm2b200_chain_device(ix->ctx, &prm->chain, n_reads, so.n_anchors, so.w.a_off, (const mm128_t*)so.w.a, st, &dev))) return rc;
	int64_t *tot = ix->h_totals.as<int64_t>() + ST_TOTALS;
	CU_TRY(cudaMemcpyAsync(tot, dev.d_totals, 64, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(ix->h_totals.p, so.w.totals, 8 * ST_TOTALS, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	const size_t tu = (size_t)tot[0], tb = (size_t)tot[1];
	if ((rc = ix->h_n_u.reserve(4 * (R + 1))) || (rc = ix->h_u_off.reserve(8 * (R + 1))) || (rc = ix->h_b_off.reserve(8 * (R + 1))) ||
	    (rc = ix->h_u.reserve(8 * (tu + 1))) || (rc = ix->h_b.reserve(16 * (tb + 1)))) return rc;
	if (R) CU_TRY(cudaMemcpyAsync(ix->h_n_u.p, dev.d_n_u, 4 * R, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(ix->h_u_off.p, dev.d_u_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(ix->h_b_off.p, dev.d_b_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	if (tu) CU_TRY(cudaMemcpyAsync(ix->h_u.p, dev.d_u, 8 * tu, cudaMemcpyDeviceToHost, st));
	if (tb) CU_TRY(cudaMemcpyAsync(ix->h_b.p, dev.d_b, 16 * tb, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	seed_finish(ix, so);
	res->chains.n_reads = n_reads;
	res->chains.n_u = ix->h_n_u.as<int32_t>(); res->chains.u_off = ix->h_u_off.as<int64_t>(); res->chains.b_off = ix->h_b_off.as<int64_t>();
	res->chains.u = ix->h_u.as<uint64_t>(); res->chains.b = ix->h_b.as<mm128_t>();
	res->chains.n_units = tot[2]; res->chains.n_sub_batches = 1; res->chains.n_visited = tot[4]; res->chains.n_lanes = 32 * tot[5];
	res->n_anchors = so.n_anchors;
	res->rep_len = ix->h_rep.as<int32_t>(); res->mini_off = ix->h_mini_off.as<int64_t>(); res->mini_pos = ix->h_mini.as<uint64_t>();
	res->a_off = ix->h_a_off.as<int64_t>(); res->a = want_anchors ? ix->h_a.as<mm128_t>() : nullptr;
	res->n_exact_sorts = ix->h_totals.as<int64_t>()[ST_EXACT];
	return MM2B200_OK;
}

namespace {
// seeding + fill + compaction; records land in ix->h_rec / h_rec_off.  Caller holds ix->mu.
int seed_fill_locked(mm2b200_index *ix, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off, const mm128_t *mv,
                     const uint32_t *bid, const int32_t *qlen, mm2b200_seed_records_t *res)
{
	int rc;
	SeedOut so;
	if ((rc = seed_stage(ix, prm, n_reads, mv_off, mv, bid, qlen, so))) return rc;
	cudaStream_t st = ix->stream;
	const size_t R = (size_t)n_reads;
	mm2b200_dev_records_t dev;
	if ((rc = mm2b200_fill_device(ix->ctx, &prm->chain, n_reads, so.n_anchors, so.w.a_off, (const mm128_t*)so.w.a, st, &dev))) return rc;
	int64_t *tot = ix->h_totals.as<int64_t>() + ST_TOTALS;
	CU_TRY(cudaMemcpyAsync(tot, dev.d_totals, 64, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	const size_t tr = (size_t)tot[3];
	if ((rc = ix->h_rec_off.reserve(8 * (R + 1))) || (rc = ix->h_rec.reserve(24 * (tr + 1)))) return rc;
	CU_TRY(cudaMemcpyAsync(ix->h_rec_off.p, dev.d_rec_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	if (tr) CU_TRY(cudaMemcpyAsync(ix->h_rec.p, dev.d_rec, 24 * tr, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	seed_finish(ix, so);
	ix->h_rec_off.as<int64_t>()[R] = (int64_t)tr;
	res->n_anchors = so.n_anchors;
	res->rec_off = ix->h_rec_off.as<int64_t>(); res->rec = (const struct new_seed*)ix->h_rec.p;
	res->rep_len = ix->h_rep.as<int32_t>(); res->mini_off = ix->h_mini_off.as<int64_t>(); res->mini_pos = ix->h_mini.as<uint64_t>();
	return MM2B200_OK;
}
} // namespace

extern "C" int mm2b200_seed_fill_batch(mm2b200_index_t *ix, const mm2b200_seed_params_t *prm, int64_t n_reads, const int64_t *mv_off,
                                       const mm128_t *mv, const uint32_t *bid, const int32_t *qlen, mm2b200_seed_records_t *res)
{
	int rc;
	if (!res) return failf(MM2B200_ERR_ARG, "NULL result");
	if ((rc = check_seed_args(ix, prm, n_reads, mv_off, mv, bid, qlen))) return rc;
	std::lock_guard<std::mutex> lk(ix->mu);
	memset(res, 0, sizeof(*res));
	return seed_fill_locked(ix, prm, n_reads, mv_off, mv, bid, qlen, res);
}

// ---------------------------------------------------------------------------------------------------------
// packets in, packets out: the device side of the driver shim (fpgadrv_shim.cu)
// ---------------------------------------------------------------------------------------------------------
// The shim's request buffers are page-locked (it hands them out itself, fpga_get_writebuf), so a request packet is DMA'd to
// HBM as it is and the minimizers are pulled out of the packet layout there; the response packets are laid out in HBM and
// DMA'd straight into page-locked response buffers.  No byte of a packet is copied by a host thread.
namespace {

__global__ void __launch_bounds__(256) k_pkt_gather(const ulonglong2 *raw, const int64_t *src16, const int64_t *mv_off, ulonglong2 *mv, int n_reads)
{
	for (int r = blockIdx.x; r < n_reads; r += gridDim.x) {
		const ulonglong2 *s = raw + src16[r];
		ulonglong2 *d = mv + mv_off[r];
		const int n = (int)(mv_off[r + 1] - mv_off[r]);
		for (int i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
	}
}

// read r's records (24 B each) and mini_pos words behind its collect_result_t, each area padded with zeros to 64 bytes
__global__ void __launch_bounds__(256) k_pkt_scatter(uint64_t *out, const int64_t *rec_dst, const int64_t *mini_dst, const int64_t *rec_off, int64_t rec_total,
                                                      const uint64_t *rec, const int64_t *mini_off, const uint64_t *mini, int n_reads)
{
	for (int r = blockIdx.x; r < n_reads; r += gridDim.x) {
		const int64_t r0 = rec_off[r], n_rec = (r + 1 < n_reads ? rec_off[r + 1] : rec_total) - r0;
		const int64_t m0 = mini_off[r], n_mini = mini_off[r + 1] - m0;
		{
			const int64_t words = 3 * n_rec, padded = (24 * n_rec + 63) / 64 * 8;
			uint64_t *d = out + rec_dst[r] / 8;
			const uint64_t *s = rec + 3 * r0;
			for (int64_t i = threadIdx.x; i < padded; i += blockDim.x) d[i] = i < words ? s[i] : 0;
		}
		{
			const int64_t padded = (8 * n_mini + 63) / 64 * 8;
			uint64_t *d = out + mini_dst[r] / 8;
			const uint64_t *s = mini + m0;
			for (int64_t i = threadIdx.x; i < padded; i += blockDim.x) d[i] = i < n_mini ? s[i] : 0;
		}
	}
}

// the 64-byte headers (chaindp_sndhdr_t, collect_result_t), built by the host
__global__ void __launch_bounds__(256) k_pkt_headers(ulonglong2 *out, const ulonglong2 *hdr, const int64_t *hdr_dst, int64_t n_hdr)
{
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < 4 * n_hdr) out[hdr_dst[i >> 2] / 16 + (i & 3)] = hdr[i];
}

} // namespace

namespace mm2b200 {

int seed_packets_begin(mm2b200_index *ix, const mm2b200_seed_params_t *prm, int n_pkts, const void *const *req, const int64_t *req_size,
                       int64_t n_reads, const int64_t *src16, const int32_t *n_mv, const uint32_t *bid, const int32_t *qlen, PacketCounts *out)
{
	int rc;
	if (!ix || !prm || !out || n_pkts <= 0 || n_reads <= 0 || n_reads >= ((int64_t)1 << 31) - 1) return failf(MM2B200_ERR_ARG, "bad packet batch");
	if (!ix->committed) return failf(MM2B200_ERR_STATE, "index not committed (mm2b200_index_commit)");
	std::lock_guard<std::mutex> lk(ix->mu);
	ix->pk_reads = -1;
	CU_TRY(cudaSetDevice(ix->device));
	cudaStream_t st = ix->stream;
	const size_t R = (size_t)n_reads;
	int64_t raw_bytes = 0, M = 0;
	for (int p = 0; p < n_pkts; ++p) raw_bytes += pkt_up64(req_size[p]);
	// per-read arrays through page-locked staging: mv_off (prefix sums), src16, bid, qlen
	if ((rc = ix->h_pkt.reserve(16 * (R + 1) + 8 * (R + 1)))) return rc;
	int64_t *h_off = ix->h_pkt.as<int64_t>(), *h_src = h_off + (R + 1);
	uint32_t *h_bid = (uint32_t*)(h_src + (R + 1));
	int32_t *h_qlen = (int32_t*)(h_bid + (R + 1));
	for (size_t r = 0; r < R; ++r) {
		if (n_mv[r] < 0 || src16[r] < 0 || 16 * (src16[r] + (int64_t)n_mv[r]) > raw_bytes) return failf(MM2B200_ERR_ARG, "read %zu lies outside its packet", r);
		h_off[r] = M; M += n_mv[r]; h_src[r] = src16[r]; h_bid[r] = bid[r]; h_qlen[r] = qlen[r];
	}
	h_off[R] = M;
	int64_t *d_off, *d_src; uint32_t *d_bid; int32_t *d_qlen;
	{
		Carver c;
		c.add(&d_off, 8 * (R + 1)); c.add(&d_src, 8 * (R + 1)); c.add(&d_bid, 4 * (R + 1)); c.add(&d_qlen, 4 * (R + 1));
		if ((rc = ix->d_pkt_meta.reserve(c.total()))) return rc;
		c.place(ix->d_pkt_meta.p);
	}
	if ((rc = ix->d_pkt_in.reserve((size_t)raw_bytes + 64)) || (rc = ix->d_pkt_mv.reserve(16 * ((size_t)M + 1)))) return rc;
	CU_TRY(cudaMemcpyAsync(d_off, h_off, 8 * (R + 1), cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_src, h_src, 8 * R, cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_bid, h_bid, 4 * R, cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_qlen, h_qlen, 4 * R, cudaMemcpyHostToDevice, st));
	const Trace tr;
	static thread_local cudaEvent_t tev[3] = { nullptr, nullptr, nullptr };
	if (tr.on) { for (auto &e : tev) if (!e) cudaEventCreate(&e); cudaEventRecord(tev[0], st); }
	int64_t at = 0;
	for (int p = 0; p < n_pkts; ++p) {
		if (req_size[p] > 0) CU_TRY(cudaMemcpyAsync((char*)ix->d_pkt_in.p + at, req[p], (size_t)req_size[p], cudaMemcpyHostToDevice, st));
		at += pkt_up64(req_size[p]);
	}
	if (tr.on) cudaEventRecord(tev[1], st);
	tr.mark("packets queued");
	const int sm = ctx_sm_count(ix->ctx);
	k_pkt_gather<<<(int)(n_reads < (int64_t)sm * 8 ? n_reads : (int64_t)sm * 8), 256, 0, st>>>((const ulonglong2*)ix->d_pkt_in.p, d_src, d_off, (ulonglong2*)ix->d_pkt_mv.p, (int)n_reads);
	CU_TRY(cudaGetLastError());
	++ix->launches;
	// collect_seed_hits, then mm_chain_dp_fpga, on the minimizers where the gather left them
	SeedOut so;
	if ((rc = seed_stage(ix, prm, n_reads, d_off, (const mm128_t*)ix->d_pkt_mv.p, d_bid, d_qlen, so, M, false))) return rc;
	mm2b200_dev_records_t dev;
	if ((rc = mm2b200_fill_device(ix->ctx, &prm->chain, n_reads, so.n_anchors, so.w.a_off, (const mm128_t*)so.w.a, st, &dev))) return rc;
	if ((rc = ix->h_pkt_cnt.reserve(8 * (R + 1) + 8 * (R + 1) + 4 * (R + 1) + 64))) return rc;
	int64_t *h_rec_off = ix->h_pkt_cnt.as<int64_t>(), *h_mini_off = h_rec_off + (R + 1), *h_tot = h_mini_off + (R + 1);
	int32_t *h_rep = (int32_t*)(h_tot + 8);
	CU_TRY(cudaMemcpyAsync(h_tot, dev.d_totals, 64, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(h_rec_off, dev.d_rec_off, 8 * R, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(h_mini_off, so.w.mini_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaMemcpyAsync(h_rep, so.w.rep_len, 4 * R, cudaMemcpyDeviceToHost, st));
	if (tr.on) cudaEventRecord(tev[2], st);
	CU_TRY(cudaStreamSynchronize(st));
	tr.mark("counts on the host");
	if (tr.on) {
		float a = 0, b = 0; cudaEventElapsedTime(&a, tev[0], tev[1]); cudaEventElapsedTime(&b, tev[1], tev[2]);
		fprintf(stderr, "[mm2b200 trace] device: %d packets (%.1f MB) uploaded in %.3f ms, gather + seeding + fill %.3f ms\n", n_pkts, 1e-6 * (double)raw_bytes, a, b);
	}
	seed_finish(ix, so);
	h_rec_off[R] = h_tot[3];
	ix->pk_rec_off = dev.d_rec_off; ix->pk_rec = dev.d_rec; ix->pk_rec_total = h_tot[3];
	ix->pk_mini_off = so.w.mini_off; ix->pk_mini = so.w.mini_out;
	ix->pk_reads = n_reads;
	out->rec_off = h_rec_off; out->mini_off = h_mini_off; out->rep_len = h_rep; out->n_anchors = so.n_anchors;
	return MM2B200_OK;
}

int seed_packets_end(mm2b200_index *ix, int n_pkts, void *const *resp, const int64_t *resp_size, int64_t n_reads,
                     const int64_t *rec_dst, const int64_t *mini_dst, int64_t n_hdr, const void *hdr, const int64_t *hdr_dst)
{
	int rc;
	if (!ix || n_pkts <= 0 || n_reads <= 0 || n_hdr <= 0) return failf(MM2B200_ERR_ARG, "bad packet batch");
	std::lock_guard<std::mutex> lk(ix->mu);
	if (ix->pk_reads != n_reads) return failf(MM2B200_ERR_STATE, "seed_packets_end without the matching seed_packets_begin");
	ix->pk_reads = -1;
	CU_TRY(cudaSetDevice(ix->device));
	cudaStream_t st = ix->stream;
	const size_t R = (size_t)n_reads, H = (size_t)n_hdr;
	int64_t out_bytes = 0;
	for (int p = 0; p < n_pkts; ++p) {
		if (resp_size[p] % 64) return failf(MM2B200_ERR_ARG, "a response packet is not a multiple of 64 bytes");
		out_bytes += resp_size[p];
	}
	if ((rc = ix->h_pkt.reserve(16 * (R + 1) + 72 * (H + 1)))) return rc;
	int64_t *h_rec_dst = ix->h_pkt.as<int64_t>(), *h_mini_dst = h_rec_dst + (R + 1), *h_hdr_dst = h_mini_dst + (R + 1);
	char *h_hdr = (char*)(h_hdr_dst + (H + 1));
	memcpy(h_rec_dst, rec_dst, 8 * R); memcpy(h_mini_dst, mini_dst, 8 * R); memcpy(h_hdr_dst, hdr_dst, 8 * H); memcpy(h_hdr, hdr, 64 * H);
	int64_t *d_rec_dst, *d_mini_dst, *d_hdr_dst; char *d_hdr;
	{
		Carver c;
		c.add(&d_rec_dst, 8 * (R + 1)); c.add(&d_mini_dst, 8 * (R + 1)); c.add(&d_hdr_dst, 8 * (H + 1)); c.add(&d_hdr, 64 * (H + 1));
		if ((rc = ix->d_pkt_meta.reserve(c.total()))) return rc;        // the arrays of _begin are no longer needed
		c.place(ix->d_pkt_meta.p);
	}
	if ((rc = ix->d_pkt_out.reserve((size_t)out_bytes + 64))) return rc;
	CU_TRY(cudaMemcpyAsync(d_rec_dst, h_rec_dst, 8 * R, cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_mini_dst, h_mini_dst, 8 * R, cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_hdr_dst, h_hdr_dst, 8 * H, cudaMemcpyHostToDevice, st));
	CU_TRY(cudaMemcpyAsync(d_hdr, h_hdr, 64 * H, cudaMemcpyHostToDevice, st));
	const int sm = ctx_sm_count(ix->ctx);
	k_pkt_scatter<<<(int)(n_reads < (int64_t)sm * 8 ? n_reads : (int64_t)sm * 8), 256, 0, st>>>((uint64_t*)ix->d_pkt_out.p, d_rec_dst, d_mini_dst, ix->pk_rec_off, ix->pk_rec_total,
	                                                                                               (const uint64_t*)ix->pk_rec, ix->pk_mini_off, ix->pk_mini, (int)n_reads);
	k_pkt_headers<<<(unsigned)((4 * H + 255) / 256), 256, 0, st>>>((ulonglong2*)ix->d_pkt_out.p, (const ulonglong2*)d_hdr, d_hdr_dst, n_hdr);
	CU_TRY(cudaGetLastError());
	ix->launches += 2;
	int64_t at = 0;
	for (int p = 0; p < n_pkts; ++p) {
		if (resp_size[p] > 0) CU_TRY(cudaMemcpyAsync(resp[p], (const char*)ix->d_pkt_out.p + at, (size_t)resp_size[p], cudaMemcpyDeviceToHost, st));
		at += resp_size[p];
	}
	CU_TRY(cudaStreamSynchronize(st));
	return MM2B200_OK;
}

} // namespace mm2b200

// ---------------------------------------------------------------------------------------------------------
// packet level: the fork's wire format
// ---------------------------------------------------------------------------------------------------------
namespace {
inline int64_t up64(int64_t x) { return pkt_up64(x); }
}

extern "C" int mm2b200_seed_work(mm2b200_index_t *ix, const mm2b200_seed_params_t *prm, const void *req, int64_t req_size,
                                 void *resp, int64_t resp_cap, int64_t *resp_size)
{
	int rc;
	if (!ix || !prm || !req || !resp || !resp_size) return failf(MM2B200_ERR_ARG, "NULL argument");
	if (!ix->committed) return failf(MM2B200_ERR_STATE, "index not committed (mm2b200_index_commit)");
	if (req_size < (int64_t)sizeof(PktHead)) return failf(MM2B200_ERR_ARG, "request packet shorter than its header");
	const char *p = (const char*)req, *end = p + req_size;
	PktHead head;
	memcpy(&head, p, sizeof(head));
	if (head.type != 3) return failf(MM2B200_ERR_ARG, "request packet type %d, expected 3 (map.c:491)", head.type);
	const int num = head.num;
	// walk the tasks (map.c:505-567)
	std::vector<PktTask> tasks((size_t)num);
	std::vector<const char*> mvp((size_t)num);
	p += sizeof(PktHead);
	int64_t n_mv = 0;
	for (int i = 0; i < num; ++i) {
		if (p + sizeof(PktTask) > end) return failf(MM2B200_ERR_ARG, "request packet truncated at task %d", i);
		memcpy(&tasks[i], p, sizeof(PktTask));
		if (tasks[i].seednum < 0) return failf(MM2B200_ERR_ARG, "task %d has a negative minimizer count", i);
		mvp[i] = p + sizeof(PktTask);
		const int64_t body = up64((int64_t)tasks[i].seednum * 16);
		if (mvp[i] + body > end && mvp[i] + (int64_t)tasks[i].seednum * 16 > end) return failf(MM2B200_ERR_ARG, "request packet truncated inside task %d", i);
		n_mv += tasks[i].seednum;
		p = mvp[i] + body;
	}
	std::lock_guard<std::mutex> lk(ix->mu);
	// reads grouped by (gap_ref, gap_qry): one device batch per group (one group in practice, map.c:358-366)
	std::vector<int> order((size_t)num), group_of((size_t)num, -1);
	std::vector<std::pair<int32_t, int32_t>> groups;
	for (int i = 0; i < num; ++i) {
		const int32_t g_ref = tasks[i].gap_ref, g_qry = tasks[i].gap_qry;
		const std::pair<int32_t, int32_t> key(g_ref, g_qry);
		size_t g = 0;
		while (g < groups.size() && groups[g] != key) ++g;
		if (g == groups.size()) groups.push_back(key);
		group_of[i] = (int)g;
	}
	struct ReadOut { std::vector<new_seed> rec; std::vector<uint64_t> mini; int32_t rep_len = 0; };
	std::vector<ReadOut> outs((size_t)num);
	std::vector<mm128_t> mv;
	std::vector<int64_t> off;
	std::vector<uint32_t> bid;
	std::vector<int32_t> qlen;
	std::vector<int> members;
	for (size_t g = 0; g < groups.size(); ++g) {
		mv.clear(); off.assign(1, 0); bid.clear(); qlen.clear(); members.clear();
		for (int i = 0; i < num; ++i) {
			if (group_of[i] != (int)g) continue;
			members.push_back(i);
			const size_t at = mv.size();
			mv.resize(at + (size_t)tasks[i].seednum);
			if (tasks[i].seednum) memcpy(mv.data() + at, mvp[i], (size_t)tasks[i].seednum * 16);
			off.push_back((int64_t)mv.size());
			bid.push_back((uint32_t)tasks[i].bid); qlen.push_back((int32_t)tasks[i].qlensum);
		}
		mm2b200_seed_params_t q = *prm;
		q.chain.max_dist_x = groups[g].first;     // max_chain_gap_ref, map.c:515
		q.chain.max_dist_y = groups[g].second;    // max_chain_gap_qry, map.c:516
		q.chain.n_segs = groups[g].second;        // map.c:517 reads n_segs from gap_qry
		mm2b200_seed_records_t rr;
		memset(&rr, 0, sizeof(rr));
		if ((rc = seed_fill_locked(ix, &q, (int64_t)members.size(), off.data(), mv.data(), bid.data(), qlen.data(), &rr))) return rc;
		for (size_t k = 0; k < members.size(); ++k) {
			ReadOut &o = outs[(size_t)members[k]];
			o.rec.assign(rr.rec + rr.rec_off[k], rr.rec + rr.rec_off[k + 1]);
			o.mini.assign(rr.mini_pos + rr.mini_off[k], rr.mini_pos + rr.mini_off[k + 1]);
			o.rep_len = rr.rep_len[k];
		}
	}
	// response packet (map.c:495-501, 533-565); padding bytes are zero here (the reference leaves them uninitialised)
	int64_t need = sizeof(PktHead);
	for (int i = 0; i < num; ++i) need += (int64_t)sizeof(PktResult) + up64((int64_t)outs[i].rec.size() * 24) + up64((int64_t)outs[i].mini.size() * 8);
	if (need > resp_cap) return failf(MM2B200_ERR_ARG, "response needs %lld bytes, buffer holds %lld", (long long)need, (long long)resp_cap);
	memset(resp, 0, (size_t)need);
	char *o = (char*)resp;
	PktHead oh;
	memset(&oh, 0, sizeof(oh));
	oh.magic = head.magic; oh.tid = head.tid; oh.num = head.num; oh.type = head.type; oh.lat = head.lat; oh.size = (uint32_t)need;
	memcpy(o, &oh, sizeof(oh));
	o += sizeof(oh);
	for (int i = 0; i < num; ++i) {
		const ReadOut &ro = outs[(size_t)i];
		const int64_t rec_b = up64((int64_t)ro.rec.size() * 24), mini_b = up64((int64_t)ro.mini.size() * 8);
		PktResult pr;
		memset(&pr, 0, sizeof(pr));
		pr.err_flag = 0; pr.read_id = tasks[i].read_id; pr.n_a = (uint32_t)ro.rec.size(); pr.n_minipos = (uint32_t)ro.mini.size();
		pr.rep_len = (uint32_t)ro.rep_len; pr.sub_size = (uint32_t)(sizeof(PktResult) + rec_b + mini_b);
		memcpy(o, &pr, sizeof(pr));
		o += sizeof(pr);
		if (!ro.rec.empty()) memcpy(o, ro.rec.data(), ro.rec.size() * 24);
		o += rec_b;
		if (!ro.mini.empty()) memcpy(o, ro.mini.data(), ro.mini.size() * 8);
		o += mini_b;
	}
	*resp_size = need;
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// bases in, regions out: sketching -> seeding -> chaining DP -> region post-processing, all in HBM (include/mm2map_b200.h)
// ---------------------------------------------------------------------------------------------------------
extern "C" int mm2b200_map_regs_batch(mm2b200_index_t *ix, const mm2b200_sketch_params_t *sk, const mm2b200_seed_params_t *prm,
                                      const mm2b200_post_params_t *post, int64_t n_reads, const int64_t *seq_off, const char *seq,
                                      const uint32_t *bid, const uint32_t *hash, mm2b200_map_result_t *res)
{
	int rc;
	if (!ix || !sk || !prm || !post || !res) return failf(MM2B200_ERR_ARG, "NULL argument");
	if (!ix->committed) return failf(MM2B200_ERR_STATE, "index not committed (mm2b200_index_commit)");
	if (n_reads < 0 || n_reads >= ((int64_t)1 << 31) - 1 || (n_reads && (!seq_off || !bid || !hash))) return failf(MM2B200_ERR_ARG, "bad batch arguments");
	if (prm->chain.n_segs > 1) return failf(MM2B200_ERR_ARG, "multi-segment reads (n_segs = %d) are outside this path", prm->chain.n_segs);
	for (int64_t r = 0; r < n_reads; ++r) {
		if (seq_off[r + 1] < seq_off[r]) return failf(MM2B200_ERR_ARG, "seq_off is not non-decreasing at read %lld", (long long)r);
		if (seq_off[r + 1] - seq_off[r] >= ((int64_t)1 << 31) - 1) return failf(MM2B200_ERR_ARG, "read %lld is longer than mm_sketch's int length", (long long)r);
	}
	const int64_t b0 = n_reads ? seq_off[0] : 0, B = n_reads ? seq_off[n_reads] - b0 : 0;
	if (B && !seq) return failf(MM2B200_ERR_ARG, "NULL sequence array");
	std::lock_guard<std::mutex> lk(ix->mu);
	const Trace tr;
	memset(res, 0, sizeof(*res));
	CU_TRY(cudaSetDevice(ix->device));
	cudaStream_t st = ix->stream;
	if (!ix->copy_stream) CU_TRY(cudaStreamCreateWithFlags(&ix->copy_stream, cudaStreamNonBlocking));
	const size_t R = (size_t)n_reads;
	// Two levels.  GROUPS of whole reads of about MM2B200_MAP_CHUNK bases (default 2 G) bound the memory: the sketching slots
	// take 16 B per base, so a group stays below 32 GB of scratch.  Each group is cut into MM2B200_MAP_STAGES chunks (default 2;
	// 1 for groups below 200 Mbases) that form a pipeline: the sketching kernels of all chunks follow the upload on their own
	// stream, and a chunk is seeded, chained and post-processed on the compute stream while the next chunk's bases are still
	// crossing PCIe and being sketched.  More chunks do not pay: every stage's tail is paid once per chunk.
	int64_t chunk_bases = 2000000000;
	{ const char *e = getenv("MM2B200_MAP_CHUNK"); const long long v = e ? atoll(e) : 0; if (v > 0) chunk_bases = (int64_t)v; }
	int n_stages = 2;
	{ const char *e = getenv("MM2B200_MAP_STAGES"); const int v = e ? atoi(e) : 0; if (v > 0 && v <= 64) n_stages = v; }
	int64_t stage_min = 200000000;                  // smaller groups are not worth a second chunk's tails (MM2B200_MAP_STAGE_MIN, bases)
	{ const char *e = getenv("MM2B200_MAP_STAGE_MIN"); const long long v = e ? atoll(e) : 0; if (v > 0) stage_min = (int64_t)v; }
	std::vector<int64_t> gcut(1, 0);
	for (int64_t r = 0; r < n_reads; ++r)
		if (seq_off[r + 1] - seq_off[gcut.back()] >= chunk_bases && r + 1 < n_reads) gcut.push_back(r + 1);
	gcut.push_back(n_reads);
	std::vector<int64_t> cut(1, 0);
	std::vector<size_t> grp_first;                  // first chunk of every group, and one behind the last
	for (size_t g = 0; g + 1 < gcut.size(); ++g) {
		grp_first.push_back(cut.size() - 1);
		const int64_t gb = n_reads ? seq_off[gcut[g + 1]] - seq_off[gcut[g]] : 0;
		const int ns = gb < stage_min ? 1 : n_stages;
		// the first chunk is the smaller one (MM2B200_MAP_FIRST percent of the group, default a third with two chunks; measured 30 / 35 / 40 / 45 %: 34.6 / 35.0 / 35.1 / 35.6 ms): what follows its
		// sketching should end when the last chunk's sketching does, which is when the upload ends
		double first = 1.0 / ns;
		if (ns == 2) first = 1.0 / 3;
		{ const char *e = getenv("MM2B200_MAP_FIRST"); const int v = e ? atoi(e) : 0; if (v > 0 && v < 100 && ns > 1) first = 0.01 * v; }
		int made = 1;
		for (int64_t r = gcut[g]; r < gcut[g + 1] && made < ns; ++r) {
			const double frac = first + (1.0 - first) * (made - 1) / (ns > 1 ? ns - 1 : 1);
			if ((double)(seq_off[r + 1] - seq_off[gcut[g]]) >= frac * (double)gb && r + 1 < gcut[g + 1]) { cut.push_back(r + 1); ++made; }
		}
		cut.push_back(gcut[g + 1]);
	}
	grp_first.push_back(cut.size() - 1);
	const size_t n_chunks = cut.size() - 1;
	char *d_seq; int64_t *d_off; uint32_t *d_bid, *d_hash; int32_t *d_qlen;
	{
		Carver c;
		c.add(&d_seq, (size_t)B + 16); c.add(&d_off, 8 * (R + n_chunks + 1)); c.add(&d_bid, 4 * (R + 1)); c.add(&d_hash, 4 * (R + 1)); c.add(&d_qlen, 4 * (R + 1));
		if ((rc = ix->d_reads.reserve(c.total()))) return rc;
		c.place(ix->d_reads.p);
	}
	if ((rc = ix->h_stage.reserve(8 * (R + n_chunks + 1) + 12 * (R + 1)))) return rc;
	int64_t *h_off = ix->h_stage.as<int64_t>();                       // chunk c's offsets, relative to its first base, at h_off + cut[c] + c
	int32_t *h_qlen = (int32_t*)(h_off + (R + n_chunks + 1));
	uint32_t *h_bid = (uint32_t*)(h_qlen + (R + 1)), *h_hash = h_bid + (R + 1);
	for (size_t c = 0; c < n_chunks; ++c)
		for (int64_t r = cut[c]; r <= cut[c + 1]; ++r) h_off[r + (int64_t)c] = seq_off[r] - seq_off[cut[c]];
	for (size_t r = 0; r < R; ++r) h_qlen[r] = (int32_t)(seq_off[r + 1] - seq_off[r]);      // qlen_sum of a single-segment read (map.c:340-341)
	// The per-read arrays go FIRST on the copy stream and from page-locked memory (bid / hash are staged: a copy from pageable
	// memory on the compute stream was seen to run only after every queued upload of bases, which held the sketching back
	// until the last base had crossed PCIe).
	if (R) { memcpy(h_bid, bid, 4 * R); memcpy(h_hash, hash, 4 * R); }
	CU_TRY(cudaMemcpyAsync(d_off, h_off, 8 * (R + n_chunks + 1), cudaMemcpyHostToDevice, ix->copy_stream));
	if (R) {
		CU_TRY(cudaMemcpyAsync(d_qlen, h_qlen, 4 * R, cudaMemcpyHostToDevice, ix->copy_stream));
		CU_TRY(cudaMemcpyAsync(d_bid, h_bid, 4 * R, cudaMemcpyHostToDevice, ix->copy_stream));
		CU_TRY(cudaMemcpyAsync(d_hash, h_hash, 4 * R, cudaMemcpyHostToDevice, ix->copy_stream));
	}
	if (!ix->ev_small) CU_TRY(cudaEventCreateWithFlags(&ix->ev_small, cudaEventDisableTiming));
	CU_TRY(cudaEventRecord(ix->ev_small, ix->copy_stream));
	CU_TRY(cudaStreamWaitEvent(st, ix->ev_small, 0));
	if (!ix->sk_stream) CU_TRY(cudaStreamCreateWithFlags(&ix->sk_stream, cudaStreamNonBlocking));
	CU_TRY(cudaStreamWaitEvent(ix->sk_stream, ix->ev_small, 0));
	while (ix->ev_sk.size() < n_chunks) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
		ix->ev_sk.push_back(e);
	}
	// every chunk's bases are queued on the copy stream now, in parts of about MM2B200_MAP_PART bases (default 48 M) with an
	// event behind each, so that the sketching of a part can start while the following parts are still on the bus
	int64_t part_bases = 48000000;
	{ const char *e = getenv("MM2B200_MAP_PART"); const long long v = e ? atoll(e) : 0; if (v > 0) part_bases = (int64_t)v; }
	std::vector<int64_t> part_cut;                 // per chunk: read boundaries of its parts, relative to the chunk's first read
	std::vector<size_t> part_first(n_chunks + 1, 0);
	for (size_t c = 0; c < n_chunks; ++c) {
		part_first[c] = part_cut.size();
		part_cut.push_back(0);
		int64_t last = cut[c];
		for (int64_t r = cut[c]; r < cut[c + 1]; ++r)
			if (seq_off[r + 1] - seq_off[last] >= part_bases && r + 1 < cut[c + 1]) { part_cut.push_back(r + 1 - cut[c]); last = r + 1; }
		part_cut.push_back(cut[c + 1] - cut[c]);
	}
	part_first[n_chunks] = part_cut.size();
	while (ix->ev_chunk.size() < part_cut.size()) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreateWithFlags(&e, tr.on ? cudaEventDefault : cudaEventDisableTiming));   // timed only for the trace
		ix->ev_chunk.push_back(e);
	}
	for (size_t c = 0; c < n_chunks; ++c) {
		const size_t p0 = part_first[c], np = part_first[c + 1] - p0 - 1;       // parts k = 0 .. np-1 use events p0 + k
		for (size_t k = 0; k < np; ++k) {
			const int64_t ra = cut[c] + part_cut[p0 + k], rb = cut[c] + part_cut[p0 + k + 1];
			const int64_t lo = seq_off[ra] - b0, nb = seq_off[rb] - seq_off[ra];
			if (nb) CU_TRY(cudaMemcpyAsync(d_seq + lo, seq + b0 + lo, (size_t)nb, cudaMemcpyHostToDevice, ix->copy_stream));
			CU_TRY(cudaEventRecord(ix->ev_chunk[p0 + k], ix->copy_stream));
		}
	}
	if ((rc = ix->h_reg_off.reserve(8 * (R + 1))) || (rc = ix->h_rep.reserve(4 * (R + 1))) || (rc = ix->h_chunk_off.reserve(8 * (R + n_chunks + 1))) ||
	    (rc = ix->h_regs.reserve(sizeof(mm2b200_reg_t) * (2 * R + 64)))) return rc;
	int64_t *h_reg_off = ix->h_reg_off.as<int64_t>();
	int64_t tot_regs = 0, tot_mv = 0, tot_anchors = 0, tot_chains = 0;
	tr.mark("uploads queued");

	size_t g = 0;                                 // the group chunk c belongs to
	for (size_t c = 0; c < n_chunks; ++c) {
		const int64_t r0 = cut[c], nr = cut[c + 1] - cut[c];
		const int64_t *d_off_c = d_off + r0 + (int64_t)c;
		while (grp_first[g + 1] <= c) ++g;
		const size_t c0 = grp_first[g], c1 = grp_first[g + 1];
		const int64_t g_r0 = cut[c0], g_b0 = n_reads ? seq_off[g_r0] : 0;
		if (c == c0) {
			// collect_minimizers (map.c:99-111) for single-segment reads: mm_sketch with rid = 0.  The kernels of every chunk of the
			// group go on the sketching stream now, part by part behind the upload; nothing of the previous group is in flight.
			const int64_t g_reads = cut[c1] - g_r0, g_bases = n_reads ? seq_off[cut[c1]] - g_b0 : 0;
			const int g_parts = (int)(part_first[c1] - part_first[c0]);
			for (size_t q = c0; q < c1; ++q) {
				const size_t p0 = part_first[q], np = part_first[q + 1] - p0 - 1;
				const int64_t q_b0 = n_reads ? seq_off[cut[q]] : 0;
				if ((rc = sketch_pipe_launch(ix->ctx, sk, q == c0, q + 1 == c1, g_reads, g_bases, g_parts, cut[q] - g_r0, q_b0 - g_b0, (int)(p0 - part_first[c0]),
				                             cut[q + 1] - cut[q], d_off + cut[q] + (int64_t)q, d_seq + (q_b0 - b0), ix->sk_stream, (int)np, part_cut.data() + p0,
				                             ix->ev_chunk.data() + p0, ix->ev_sk[q]))) return rc;
			}
			tr.mark("sketching queued");
		}
		// the chunk's minimizers: counted and packed once its parts are sketched
		mm2b200_dev_sketch_t dsk;
		int64_t n_mv = 0;
		if ((rc = sketch_pipe_finish(ix->ctx, r0 - g_r0, (n_reads ? seq_off[r0] : 0) - g_b0, nr, d_off_c, st, ix->ev_sk[c], &dsk, &n_mv))) return rc;
		tr.mark("sketch counted");
		// collect_seed_hits (map.c:187-236) on the minimizers where the sketch left them
		SeedOut so;
		if ((rc = seed_stage(ix, prm, nr, dsk.d_mv_off, dsk.d_mv, d_bid + r0, d_qlen + r0, so, n_mv, false))) return rc;
		tr.mark("seeding queued");
		// mm_chain_dp, then mm_gen_regs + chain_post (map.c:837-866)
		mm2b200_dev_result_t dev;
		if ((rc = mm2b200_chain_device(ix->ctx, &prm->chain, nr, so.n_anchors, so.w.a_off, (const mm128_t*)so.w.a, st, &dev))) return rc;
		const int64_t cap = so.n_anchors / (prm->chain.min_cnt > 1 ? prm->chain.min_cnt : 1) + 1;
		mm2b200_dev_regs_t dr;
		if ((rc = mm2b200_post_device(ix->ctx, post, nr, &dev, cap, d_hash + r0, d_qlen + r0, st, &dr))) return rc;
		int64_t *tot = ix->h_totals.as<int64_t>() + ST_TOTALS;
		int64_t *h_co = ix->h_chunk_off.as<int64_t>() + r0 + (int64_t)c;
		CU_TRY(cudaMemcpyAsync(tot, dr.d_totals, 8, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaMemcpyAsync(tot + 1, dev.d_totals, 16, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaMemcpyAsync(h_co, dr.d_reg_off, 8 * ((size_t)nr + 1), cudaMemcpyDeviceToHost, st));
		if (nr) CU_TRY(cudaMemcpyAsync(ix->h_rep.as<int32_t>() + r0, so.w.rep_len, 4 * (size_t)nr, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaStreamSynchronize(st));
		tr.mark("chains + regions done");
		const int64_t n_regs = tot[0];
		if ((rc = ix->h_regs.reserve_keep(sizeof(mm2b200_reg_t) * (size_t)(tot_regs + n_regs + 1), sizeof(mm2b200_reg_t) * (size_t)tot_regs))) return rc;
		if (n_regs) CU_TRY(cudaMemcpyAsync((mm2b200_reg_t*)ix->h_regs.p + tot_regs, dr.d_regs, sizeof(mm2b200_reg_t) * (size_t)n_regs, cudaMemcpyDeviceToHost, st));
		for (int64_t r = 0; r < nr; ++r) h_reg_off[r0 + r] = tot_regs + h_co[r];
		seed_finish(ix, so);
		tot_regs += n_regs; tot_mv += n_mv; tot_anchors += so.n_anchors; tot_chains += tot[1];
		{                                          // the regions must be off the device before the next chunk reuses the buffers
			CU_TRY(cudaStreamSynchronize(st));
			if (c + 1 == c1) { CU_TRY(cudaStreamSynchronize(ix->sk_stream)); sketch_stage_account(ix->ctx); }
			post_stage_account(ix->ctx);
		}
	}
	h_reg_off[R] = tot_regs;
	CU_TRY(cudaStreamSynchronize(ix->copy_stream));
	tr.mark("regions on the host");
	res->n_reads = n_reads; res->reg_off = ix->h_reg_off.as<int64_t>(); res->regs = (const mm2b200_reg_t*)ix->h_regs.p;
	res->rep_len = ix->h_rep.as<int32_t>();
	res->n_minimizers = tot_mv; res->n_anchors = tot_anchors; res->n_chains = tot_chains;
	return MM2B200_OK;
}
