// chain_api.cu -- host side of libmm2chain_b200.so: contexts, the sub-batch pipeline (H2D / kernels / D2H on
// per-slot streams), the reference entry points and their combining batcher.  C ABI in include/mm2chain_b200.h.
//
// There is deliberately no CPU implementation of any stage in this file: if CUDA is unavailable every compute
// entry point fails with MM2B200_ERR_CUDA.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "chain_internal.h"

using namespace mm2b200;

// kalloc of the host program (kalloc.c:72,120), if it exports it
extern "C" void *kmalloc(void *km, size_t size) __attribute__((weak));
extern "C" void kfree(void *km, void *ptr) __attribute__((weak));

static thread_local std::string g_err;

// Long-read batches keep up to 32 sub-batches in flight on separate streams; with the default 8 hardware queues
// streams share queues and serialise behind each other's long fill kernels (measured: 1021 -> 540 ms on the 6000-read
// map-pb slice).  The variable is read when the CUDA context is created, so it is set when the library is loaded and
// never overrides a value the user chose.
__attribute__((constructor)) static void mm2b200_on_load() { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); }

static int fail(int code, const char *fmt, ...)
{
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_err = buf;
	return code;
}

#define CUDA_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
	return fail(MM2B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------------------------------------------------
// growable pinned host buffer (the context's "kalloc-style" arena for results)
// ---------------------------------------------------------------------------------------------------------
struct PinnedBuf {
	void *ptr = nullptr;
	size_t cap = 0;
	int reserve(size_t bytes, size_t keep = 0)
	{
		if (bytes <= cap) return MM2B200_OK;
		size_t ncap = cap ? cap : (size_t)1 << 16;
		while (ncap < bytes) ncap += ncap / 2 + 4096;
		void *np = nullptr;
		if (cudaHostAlloc(&np, ncap, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return fail(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu) failed", ncap); }
		if (keep && ptr) memcpy(np, ptr, keep);
		if (ptr) cudaFreeHost(ptr);
		ptr = np; cap = ncap;
		return MM2B200_OK;
	}
	void release() { if (ptr) cudaFreeHost(ptr); ptr = nullptr; cap = 0; }
	template <class T> T *as() { return (T*)ptr; }
};

struct StageAcc { double ms = 0; int64_t launches = 0; };
struct StageSpan { int stage; cudaEvent_t beg, end; int launches; };
static const char *const STAGE_NAMES[] = { "prep", "fill", "compact", "tail", "gather" };
enum { ST_PREP = 0, ST_FILL, ST_COMPACT, ST_TAIL, ST_GATHER, ST_N };

// ---------------------------------------------------------------------------------------------------------
// host batcher: worker threads that encode a sub-batch of anchors into the 8-byte wire format while the previous
// sub-batch is on the PCIe bus (the upload is the end-to-end bottleneck: 16 B/anchor at ~55 GB/s)
// ---------------------------------------------------------------------------------------------------------
struct WorkerPool {
	std::vector<std::thread> th;
	std::mutex mu;
	std::condition_variable cv_go, cv_done;
	uint64_t gen = 0;
	int pending = 0;
	bool stop = false;
	const std::function<void(int)> *fn = nullptr;
	int size() const { return (int)th.size() + 1; }       // the caller is worker 0
	void start(int n)
	{
		uint64_t g0;
		{ std::lock_guard<std::mutex> lk(mu); g0 = gen; }
		for (int k = 1; k < n; ++k) th.emplace_back([this, k, g0] { loop(k, g0); });
	}
	void loop(int k, uint64_t seen)
	{
		for (;;) {
			std::unique_lock<std::mutex> lk(mu);
			cv_go.wait(lk, [&] { return stop || gen != seen; });
			if (stop) return;
			seen = gen;
			const std::function<void(int)> *f = fn;
			lk.unlock();
			(*f)(k);
			lk.lock();
			if (--pending == 0) cv_done.notify_one();
		}
	}
	void run(const std::function<void(int)> &f)
	{
		if (!th.empty()) {
			{ std::lock_guard<std::mutex> lk(mu); fn = &f; pending = (int)th.size(); ++gen; }
			cv_go.notify_all();
		}
		f(0);
		if (!th.empty()) {
			std::unique_lock<std::mutex> lk(mu);
			cv_done.wait(lk, [&] { return pending == 0; });
		}
	}
	void shutdown()
	{
		{ std::lock_guard<std::mutex> lk(mu); stop = true; }
		cv_go.notify_all();
		for (auto &t : th) t.join();
		th.clear();
		stop = false;
	}
};

struct WirePart {                 // what one worker produced for its range of blocks
	std::vector<uint32_t> dict;   // dictionaries of its blocks, back to back
	bool ok = true;
};

// Encode blocks [b0, b1) of a sub-batch (format: chain_kernels.cu, k_unpack_wire).  tab[b] receives the block's
// descriptor with a dictionary position local to this part; the caller rebases it.  Returns false when a block does
// not fit the format (its qpos values need more bits than the block's flags|qspan dictionary leaves free).
static bool wire_encode(const mm128_t *src, int64_t n, int64_t b0, int64_t b1, uint64_t *out, uint64_t *tab, std::vector<uint32_t> &dict)
{
	dict.clear();
	for (int64_t b = b0; b < b1; ++b) {
		const int64_t k0 = b << WIRE_BLOCK_SHIFT, k1 = k0 + WIRE_BLOCK < n ? k0 + WIRE_BLOCK : n;
		uint32_t hx[WIRE_BLOCK + 1], hy[WIRE_BLOCK], lastx, qor = 0, ydiff = 0;
		int nhx = 0, nhy = 1, ybits = 0;
		// Common case in one pass: flags|seg|qspan is the same for the whole block (no y index bits).  The x dictionary gets
		// one entry per change of strand|rid -- x is sorted inside a read, so a change is almost always a new value and
		// looking for an earlier copy is not worth it; 256 anchors can never need more than 256 entries.
		hy[0] = (uint32_t)(src[k0].y >> 32);
		lastx = ~(uint32_t)(src[k0].x >> 32);
		for (int64_t k = k0; k < k1; ++k) {
			const uint64_t x = src[k].x, y = src[k].y;
			const uint32_t vx = (uint32_t)(x >> 32);
			hx[nhx] = vx;                                 // branch-free: the slot is claimed only when the value changed
			nhx += vx != lastx;
			lastx = vx;
			ydiff |= (uint32_t)(y >> 32) ^ hy[0];
			qor |= (uint32_t)y;
			out[k] = (x & 0xffffffffull) | (y << 32) | ((uint64_t)(nhx - 1) << 56);
		}
		if (ydiff) {                                      // several flags|seg|qspan values: second pass with a y index per anchor
			uint8_t iy[WIRE_BLOCK], slot[256];
			memset(slot, 0, sizeof(slot));
			uint32_t lasty = hy[0];
			int cy = 0;
			for (int64_t k = k0; k < k1; ++k) {
				const uint32_t vy = (uint32_t)(src[k].y >> 32);
				if (vy != lasty) {
					cy = slot[vy & 0xff];                     // direct-mapped on qspan, the part that varies (HPC minimizers)
					if (hy[cy] != vy) {
						for (cy = 0; cy < nhy && hy[cy] != vy; ++cy) {}
						if (cy == nhy) hy[nhy++] = vy;
						slot[vy & 0xff] = (uint8_t)cy;
					}
					lasty = vy;
				}
				iy[k - k0] = (uint8_t)cy;
			}
			while ((1 << ybits) < nhy) ++ybits;
			if (qor >> (24 - ybits)) return false;
			for (int64_t k = k0; k < k1; ++k) out[k] |= (uint64_t)iy[k - k0] << (32 + 24 - ybits);
		} else if (qor >> 24) return false;               // a qpos of 2^24 or more
		tab[b] = (uint64_t)dict.size() | ((uint64_t)nhx << 32) | ((uint64_t)ybits << 48);
		dict.insert(dict.end(), hx, hx + nhx);
		dict.insert(dict.end(), hy, hy + nhy);
	}
	return true;
}

// ---------------------------------------------------------------------------------------------------------
// one in-flight sub-batch
// ---------------------------------------------------------------------------------------------------------
struct Slot {
	cudaStream_t stream = nullptr;
	cudaEvent_t ev_totals = nullptr, ev_done = nullptr;
	// mm2b200_chain_device enqueues on the CALLER's stream and returns: the workspace (and the device result pointers
	// handed out) stay in use until ev_user fires.  Anything that reuses or frees the slot from another stream waits for it.
	cudaEvent_t ev_user = nullptr;
	cudaStream_t user_stream = nullptr;
	bool user_pending = false;
	Workspace w{};
	void *d_block = nullptr;          // one allocation carved into the workspace arrays
	void *d_rec_out = nullptr;        // lazily allocated (fill_batch only)
	size_t cap_anchors = 0, cap_reads = 0, cap_rec_out = 0;
	PinnedBuf h_a, h_offs, h_totals;
	// the chunk currently occupying the slot
	int state = 0;                    // 0 idle, 1 kernels enqueued (totals pending), 2 outputs in flight
	int64_t read0 = 0, n_reads = 0, anchor0 = 0, n_anchors = 0;
	int64_t ubase = 0, bbase = 0;     // position of this chunk's outputs in the packed result
};

struct mm2b200_ctx {
	int device = 0;
	int profile = 0;
	int64_t slot_anchors = 0;
	bool adaptive_slots = true;
	LaunchCfg lc{};
	std::vector<Slot> slots;
	std::mutex mu;
	std::atomic<int64_t> launches{0};
	// upload wire format (host batcher)
	bool wire = false;
	WorkerPool pool;
	std::vector<WirePart> parts;
	int64_t h2d_bytes = 0, wire_sub_batches = 0;   // of the last batch call
	// results (pinned)
	PinnedBuf r_n_u, r_u_off, r_b_off, r_u, r_b, r_rec_off, r_rec;
	// mm2b200_chain_device runs the two halves of a batch on the caller's stream and on this one
	cudaStream_t aux_stream = nullptr;
	cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
	bool split_device = true;
	// state of the stages built around the DP (post-processing, sketching): owned by their own translation units
	void *ext[4] = { nullptr, nullptr, nullptr, nullptr };
	void (*ext_free[4])(void*) = { nullptr, nullptr, nullptr, nullptr };
	// profiling
	StageAcc acc[ST_N];
	std::vector<StageSpan> spans;
	std::vector<cudaEvent_t> ev_pool;
};

static int ensure_slot(mm2b200_ctx *ctx, Slot &s, size_t anchors, size_t reads)
{
	if (anchors <= s.cap_anchors && reads <= s.cap_reads && s.d_block) return MM2B200_OK;
	size_t A = s.cap_anchors > anchors ? s.cap_anchors : anchors;
	size_t R = s.cap_reads > reads ? s.cap_reads : reads;
	if (A < 1024) A = 1024;
	if (R < 64) R = 64;
	if (A >= ((size_t)1 << 31) - 64) return fail(MM2B200_ERR_ARG, "sub-batch of %zu anchors exceeds the 2^31 index space", A);
	CUDA_TRY(cudaStreamSynchronize(s.stream));
	if (s.user_pending) { CUDA_TRY(cudaEventSynchronize(s.ev_user)); s.user_pending = false; }
	if (s.d_block) { cudaFree(s.d_block); s.d_block = nullptr; }
	struct Item { void **pp; size_t bytes; };
	Workspace &w = s.w;
	void *pa = nullptr, *poffs = nullptr;
	Item items[] = {
		{ &pa, 16 * A }, { &poffs, 8 * (R + 1) },
		{ (void**)&w.avg, 4 * R }, { (void**)&w.units, 2 * 16 * (A / 2 + 1) }, { (void**)&w.counters, 2 * 4 * CNT_N }, { (void**)&w.visited, 16 },
		{ (void**)&w.wrec, 16 * A }, { (void**)&w.f, 4 * A }, { (void**)&w.p, 4 * A }, { (void**)&w.v, 4 * A }, { (void**)&w.t, 4 * A },
		{ (void**)&w.newid, 4 * A }, { (void**)&w.rec, 24 * A }, { (void**)&w.recp, 4 * A }, { (void**)&w.new_i, 4 * R },
		{ (void**)&w.succ, A }, { (void**)&w.claimed, A }, { (void**)&w.members, 4 * A },
		{ (void**)&w.cand, 8 * A }, { (void**)&w.chain_u, 8 * A }, { (void**)&w.wy, 8 * A },
		{ (void**)&w.voff, 4 * A }, { (void**)&w.order, 4 * A }, { (void**)&w.ooff, 4 * A }, { (void**)&w.stk, 4 * A },
		{ (void**)&w.n_u, 4 * R }, { (void**)&w.n_v, 4 * R },
		{ (void**)&w.u_off, 8 * (R + 1) }, { (void**)&w.b_off, 8 * (R + 1) }, { (void**)&w.rec_off, 8 * (R + 1) },
		{ (void**)&w.totals, 8 * 8 }, { (void**)&w.scan_part, 8 * 3 * (R / 1024 + 2) }, { (void**)&w.u_out, 8 * A }, { (void**)&w.b_out, 16 * A },
	};
	size_t total = 0;
	for (auto &it : items) total += align_up(it.bytes, 256);
	if (cudaMalloc(&s.d_block, total) != cudaSuccess) {
		cudaGetLastError();
		s.cap_anchors = s.cap_reads = 0;
		return fail(MM2B200_ERR_NOMEM, "cudaMalloc(%zu bytes) for a %zu-anchor sub-batch failed", total, A);
	}
	char *cur = (char*)s.d_block;
	for (auto &it : items) { *it.pp = cur; cur += align_up(it.bytes, 256); }
	w.a = (const ulonglong2*)pa;
	w.offs = (const int64_t*)poffs;
	w.units_cap = (int32_t)(A / 2 + 1);
	s.cap_anchors = A; s.cap_reads = R;
	(void)ctx;
	return MM2B200_OK;
}

static int ensure_rec_out(Slot &s)
{
	if (s.cap_rec_out >= s.cap_anchors && s.d_rec_out) return MM2B200_OK;
	if (s.d_rec_out) cudaFree(s.d_rec_out);
	s.d_rec_out = nullptr;
	if (cudaMalloc(&s.d_rec_out, 24 * s.cap_anchors) != cudaSuccess) { cudaGetLastError(); return fail(MM2B200_ERR_NOMEM, "cudaMalloc(rec_out) failed"); }
	s.cap_rec_out = s.cap_anchors;
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// profiling helpers
// ---------------------------------------------------------------------------------------------------------
static cudaEvent_t get_event(mm2b200_ctx *ctx)
{
	cudaEvent_t e = nullptr;
	if (!ctx->ev_pool.empty()) { e = ctx->ev_pool.back(); ctx->ev_pool.pop_back(); }
	else cudaEventCreate(&e);
	return e;
}

struct StageTimer {
	mm2b200_ctx *ctx; cudaStream_t st; int stage; cudaEvent_t beg = nullptr;
	StageTimer(mm2b200_ctx *c, cudaStream_t s, int stage_) : ctx(c), st(s), stage(stage_)
	{
		if (ctx->profile) { beg = get_event(ctx); cudaEventRecord(beg, st); }
	}
	void done(int launches)
	{
		ctx->launches += launches;
		if (ctx->profile) {
			cudaEvent_t end = get_event(ctx);
			cudaEventRecord(end, st);
			ctx->spans.push_back(StageSpan{stage, beg, end, launches});
		} else ctx->acc[stage].launches += launches;
	}
};

static void drain_spans(mm2b200_ctx *ctx)
{
	for (auto &sp : ctx->spans) {
		float ms = 0;
		cudaEventSynchronize(sp.end);
		cudaEventElapsedTime(&ms, sp.beg, sp.end);
		ctx->acc[sp.stage].ms += ms;
		ctx->acc[sp.stage].launches += sp.launches;
		ctx->ev_pool.push_back(sp.beg);
		ctx->ev_pool.push_back(sp.end);
	}
	ctx->spans.clear();
}

// ---------------------------------------------------------------------------------------------------------
// kernel sequences
// ---------------------------------------------------------------------------------------------------------
static int check_params(const mm2b200_params_t *p)
{
	if (!p) return fail(MM2B200_ERR_ARG, "params is NULL");
	if (p->max_dist_x < 0 || p->max_dist_y < 0) return fail(MM2B200_ERR_ARG, "max_dist_x/max_dist_y must be >= 0");
	if (p->max_iter < 0) return fail(MM2B200_ERR_ARG, "max_iter must be >= 0 (0 = unbounded)");
	return MM2B200_OK;
}

// what the kernels get: max_skip beyond 2^30 behaves like 2^30 (a read has fewer than 2^29 anchors, so the budget can
// never run out) and keeps max_skip + 1 inside int32
static mm2b200_params_t kernel_params(const mm2b200_params_t &p)
{
	mm2b200_params_t q = p;
	if (q.max_skip > (1 << 30)) q.max_skip = 1 << 30;
	return q;
}

// prep + fill + compaction (the mm_chain_dp_fpga half, chain.c:218-327)
static int enqueue_fill(mm2b200_ctx *ctx, const Workspace &w, const mm2b200_params_t &prm, cudaStream_t st)
{
	{ StageTimer t(ctx, st, ST_PREP); t.done(launch_prep(w, prm, st)); }
	{ StageTimer t(ctx, st, ST_FILL); t.done(launch_fill(w, kernel_params(prm), ctx->lc, prm.is_cdna != 0, st)); }
	return MM2B200_OK;
}
static int enqueue_compact(mm2b200_ctx *ctx, const Workspace &w, const mm2b200_params_t &prm, cudaStream_t st)
{
	StageTimer t(ctx, st, ST_COMPACT); t.done(launch_compact(w, prm, st));
	return MM2B200_OK;
}
// ends / backtrack / order / gather (the mm_chain_dp_bottom half, chain.c:329-431)
static int enqueue_bottom(mm2b200_ctx *ctx, const Workspace &w, const mm2b200_params_t &prm, cudaStream_t st)
{
	{ StageTimer t(ctx, st, ST_TAIL); t.done(launch_tail(w, prm, st)); }
	{ StageTimer t(ctx, st, ST_GATHER); t.done(launch_gather(w, st)); }
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------------------
extern "C" const char *mm2b200_last_error(void) { return g_err.c_str(); }
extern "C" const char *mm2b200_version(void) { return "mm2chain-b200 0.1 (sm_100a)"; }

extern "C" int mm2b200_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

extern "C" void *mm2b200_pinned_alloc(size_t bytes)
{
	void *p = nullptr;
	if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); fail(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu) failed", bytes); return nullptr; }
	return p;
}
extern "C" void mm2b200_pinned_free(void *p) { if (p) cudaFreeHost(p); }

static int add_slot(mm2b200_ctx *ctx)
{
	ctx->slots.emplace_back();
	Slot &s = ctx->slots.back();
	if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaEventCreateWithFlags(&s.ev_totals, cudaEventDisableTiming) != cudaSuccess ||
	    cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming) != cudaSuccess ||
	    cudaEventCreateWithFlags(&s.ev_user, cudaEventDisableTiming) != cudaSuccess)
		return fail(MM2B200_ERR_CUDA, "stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
	return MM2B200_OK;
}

extern "C" int mm2b200_create(const mm2b200_config_t *cfg, mm2b200_ctx_t **out)
{
	if (!out) return fail(MM2B200_ERR_ARG, "ctx out pointer is NULL");
	*out = nullptr;
	mm2b200_config_t c{};
	if (cfg) c = *cfg;
	int ndev = 0;
	CUDA_TRY(cudaGetDeviceCount(&ndev));
	if (c.device < 0 || c.device >= ndev) return fail(MM2B200_ERR_CUDA, "device %d not available (%d CUDA devices)", c.device, ndev);
	CUDA_TRY(cudaSetDevice(c.device));
	mm2b200_ctx *ctx = new mm2b200_ctx();
	ctx->device = c.device;
	ctx->profile = c.profile;
	ctx->slot_anchors = c.slot_anchors > 0 ? c.slot_anchors : ((int64_t)8 << 20);
	const int ns = c.n_slots > 0 ? c.n_slots : 3;
	ctx->adaptive_slots = c.n_slots <= 0;      // default: long-read batches may add slots (see run_batch)
	for (int k = 0; k < ns; ++k) {
		int rc = add_slot(ctx);
		if (rc) { mm2b200_destroy(ctx); return rc; }
	}
	query_launch_cfg(&ctx->lc);
	ctx->lc.count_visited = c.count_visited;
	if (cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
	    cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) != cudaSuccess) {
		int rc = fail(MM2B200_ERR_CUDA, "stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
		mm2b200_destroy(ctx);
		return rc;
	}
	{ const char *es = getenv("MM2B200_SPLIT"); ctx->split_device = !(es && atoi(es) == 0); }
	{	// host batcher threads: opt-in (MM2B200_WIRE=1 or mm2b200_set_upload; MM2B200_WIRE_THREADS overrides the thread count).
		// Off by default: on the measured GPU box (16 vCPUs) encoding 77 M anchors costs more host memory time than the
		// 0.55 GB it keeps off the PCIe bus (e2e 41 ms vs 27 ms raw, profiles/r01j_wire_upload.md).
		const char *ew = getenv("MM2B200_WIRE"), *et = getenv("MM2B200_WIRE_THREADS");
		ctx->wire = ew && atoi(ew) != 0;
		int nt = et ? atoi(et) : (int)std::thread::hardware_concurrency();
		if (nt > 16) nt = 16;
		if (nt < 1) nt = 1;
		if (ctx->wire) ctx->pool.start(nt);
	}
	*out = ctx;
	return MM2B200_OK;
}

extern "C" void mm2b200_destroy(mm2b200_ctx_t *ctx)
{
	if (!ctx) return;
	ctx->pool.shutdown();
	cudaSetDevice(ctx->device);
	for (auto &s : ctx->slots) {
		if (s.stream) cudaStreamSynchronize(s.stream);
		if (s.d_block) cudaFree(s.d_block);
		if (s.d_rec_out) cudaFree(s.d_rec_out);
		s.h_a.release(); s.h_offs.release(); s.h_totals.release();
		if (s.ev_totals) cudaEventDestroy(s.ev_totals);
		if (s.ev_done) cudaEventDestroy(s.ev_done);
		if (s.ev_user) cudaEventDestroy(s.ev_user);
		if (s.stream) cudaStreamDestroy(s.stream);
	}
	for (int k = 0; k < 4; ++k) if (ctx->ext[k] && ctx->ext_free[k]) { ctx->ext_free[k](ctx->ext[k]); ctx->ext[k] = nullptr; }
	drain_spans(ctx);
	if (ctx->aux_stream) { cudaStreamSynchronize(ctx->aux_stream); cudaStreamDestroy(ctx->aux_stream); }
	if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
	if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
	for (auto e : ctx->ev_pool) cudaEventDestroy(e);
	ctx->r_n_u.release(); ctx->r_u_off.release(); ctx->r_b_off.release(); ctx->r_u.release(); ctx->r_b.release();
	ctx->r_rec_off.release(); ctx->r_rec.release();
	delete ctx;
}

extern "C" int64_t mm2b200_launch_count(const mm2b200_ctx_t *ctx) { return ctx ? ctx->launches.load() : 0; }

extern "C" int mm2b200_set_split(mm2b200_ctx_t *ctx, int32_t on)
{
	if (!ctx) return fail(MM2B200_ERR_ARG, "NULL context");
	std::lock_guard<std::mutex> lk(ctx->mu);
	ctx->split_device = on != 0;
	return MM2B200_OK;
}

extern "C" int mm2b200_set_upload(mm2b200_ctx_t *ctx, int32_t wire, int32_t threads)
{
	if (!ctx || threads < 0) return fail(MM2B200_ERR_ARG, "bad upload settings");
	std::lock_guard<std::mutex> lk(ctx->mu);
	ctx->pool.shutdown();
	ctx->wire = wire != 0;
	int nt = threads ? threads : (int)std::thread::hardware_concurrency();
	if (nt > 64) nt = 64;
	if (nt < 1) nt = 1;
	if (ctx->wire) ctx->pool.start(nt);
	return MM2B200_OK;
}

extern "C" int mm2b200_wire_encode(const mm128_t *a, int64_t n, void *buf, size_t cap, size_t *bytes, size_t *tab_pos, size_t *dict_pos)
{
	if (n < 0 || (n > 0 && !a) || !buf || !bytes || !tab_pos || !dict_pos) return fail(MM2B200_ERR_ARG, "bad wire_encode arguments");
	const int64_t nblk = (n + WIRE_BLOCK - 1) >> WIRE_BLOCK_SHIFT;
	const size_t pos_tab = (size_t)n * 8, pos_dict = pos_tab + (size_t)nblk * 8;
	if (cap < pos_dict) return fail(MM2B200_ERR_ARG, "wire buffer too small (%zu < %zu)", cap, pos_dict);
	std::vector<uint32_t> dict;
	if (!wire_encode(a, n, 0, nblk, (uint64_t*)buf, (uint64_t*)((char*)buf + pos_tab), dict)) return 1;
	if (cap < pos_dict + dict.size() * 4) return fail(MM2B200_ERR_ARG, "wire buffer too small for the dictionaries");
	if (!dict.empty()) memcpy((char*)buf + pos_dict, dict.data(), dict.size() * 4);
	*bytes = pos_dict + dict.size() * 4; *tab_pos = pos_tab; *dict_pos = pos_dict;
	return MM2B200_OK;
}

extern "C" int mm2b200_upload_stats(const mm2b200_ctx_t *ctx, int64_t *h2d_bytes, int64_t *wire_sub_batches)
{
	if (!ctx) return fail(MM2B200_ERR_ARG, "NULL ctx");
	if (h2d_bytes) *h2d_bytes = ctx->h2d_bytes;
	if (wire_sub_batches) *wire_sub_batches = ctx->wire_sub_batches;
	return MM2B200_OK;
}

extern "C" int mm2b200_stage_ms(mm2b200_ctx_t *ctx, const char *stage, double *ms, int64_t *launches)
{
	if (!ctx || !stage) return fail(MM2B200_ERR_ARG, "NULL argument");
	std::lock_guard<std::mutex> lk(ctx->mu);
	cudaSetDevice(ctx->device);
	drain_spans(ctx);
	for (int i = 0; i < ST_N; ++i)
		if (!strcmp(stage, STAGE_NAMES[i])) {
			if (ms) *ms = ctx->acc[i].ms;
			if (launches) *launches = ctx->acc[i].launches;
			return MM2B200_OK;
		}
	return fail(MM2B200_ERR_ARG, "unknown stage '%s'", stage);
}

extern "C" void mm2b200_stage_reset(mm2b200_ctx_t *ctx)
{
	if (!ctx) return;
	std::lock_guard<std::mutex> lk(ctx->mu);
	cudaSetDevice(ctx->device);
	drain_spans(ctx);
	for (int i = 0; i < ST_N; ++i) ctx->acc[i] = StageAcc();
}

// ---------------------------------------------------------------------------------------------------------
// sub-batch pipeline
// ---------------------------------------------------------------------------------------------------------
enum Mode { MODE_CHAIN = 0, MODE_FILL = 1, MODE_BOTTOM = 2 };

struct BatchJob {
	mm2b200_ctx *ctx;
	const mm2b200_params_t *prm;
	Mode mode;
	int64_t n_reads;
	const int64_t *offs;       // host, [n_reads+1]
	const char *src;           // host anchors (16 B) or records (24 B)
	size_t elem;               // 16 or 24
	bool src_pinned;
	int32_t *dbg_f, *dbg_p, *dbg_v;   // fill mode only
	int64_t ubase = 0, bbase = 0;     // running output positions
	int64_t n_units = 0, n_chunks = 0, n_visited = 0, n_lanes = 0;
};

static bool is_pinned(const void *p)
{
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
	return at.type == cudaMemoryTypeHost;
}

// stage A: totals are on the host -> enqueue the output copies
static int finish_a(BatchJob &job, Slot &s)
{
	mm2b200_ctx *ctx = job.ctx;
	CUDA_TRY(cudaEventSynchronize(s.ev_totals));
	const int64_t *tot = s.h_totals.as<int64_t>();
	const int64_t tu = tot[0], tb = tot[1];
	job.n_units += tot[2];
	job.n_visited += tot[4];
	job.n_lanes += 32 * tot[5];
	s.ubase = job.ubase; s.bbase = job.bbase;
	// (result arenas were reserved to their upper bounds in run_batch, so nothing here can move a buffer that an
	// earlier chunk's copy is still writing to; per-chunk offset copies cover n_reads entries so chunks never overlap)
	if (job.mode == MODE_FILL) {
		const int64_t tr = tot[3];
		CUDA_TRY(cudaMemcpyAsync(ctx->r_rec_off.as<int64_t>() + s.read0, s.w.rec_off, (size_t)s.n_reads * 8, cudaMemcpyDeviceToHost, s.stream));
		if (tr) CUDA_TRY(cudaMemcpyAsync(ctx->r_rec.as<char>() + (size_t)job.bbase * 24, s.w.rec_out, (size_t)tr * 24, cudaMemcpyDeviceToHost, s.stream));
		if (job.dbg_f) CUDA_TRY(cudaMemcpyAsync(job.dbg_f + s.anchor0, s.w.f, (size_t)s.n_anchors * 4, cudaMemcpyDeviceToHost, s.stream));
		if (job.dbg_p) CUDA_TRY(cudaMemcpyAsync(job.dbg_p + s.anchor0, s.w.p, (size_t)s.n_anchors * 4, cudaMemcpyDeviceToHost, s.stream));
		if (job.dbg_v) CUDA_TRY(cudaMemcpyAsync(job.dbg_v + s.anchor0, s.w.v, (size_t)s.n_anchors * 4, cudaMemcpyDeviceToHost, s.stream));
		job.bbase += tr;
	} else {
		CUDA_TRY(cudaMemcpyAsync(ctx->r_n_u.as<int32_t>() + s.read0, s.w.n_u, (size_t)s.n_reads * 4, cudaMemcpyDeviceToHost, s.stream));
		CUDA_TRY(cudaMemcpyAsync(ctx->r_u_off.as<int64_t>() + s.read0, s.w.u_off, (size_t)s.n_reads * 8, cudaMemcpyDeviceToHost, s.stream));
		CUDA_TRY(cudaMemcpyAsync(ctx->r_b_off.as<int64_t>() + s.read0, s.w.b_off, (size_t)s.n_reads * 8, cudaMemcpyDeviceToHost, s.stream));
		if (tu) CUDA_TRY(cudaMemcpyAsync(ctx->r_u.as<uint64_t>() + job.ubase, s.w.u_out, (size_t)tu * 8, cudaMemcpyDeviceToHost, s.stream));
		if (tb) CUDA_TRY(cudaMemcpyAsync(ctx->r_b.as<char>() + (size_t)job.bbase * 16, s.w.b_out, (size_t)tb * 16, cudaMemcpyDeviceToHost, s.stream));
		job.ubase += tu; job.bbase += tb;
	}
	CUDA_TRY(cudaEventRecord(s.ev_done, s.stream));
	s.state = 2;
	return MM2B200_OK;
}

// stage B: outputs are on the host -> rebase the chunk-local offsets
static int finish_b(BatchJob &job, Slot &s)
{
	mm2b200_ctx *ctx = job.ctx;
	CUDA_TRY(cudaEventSynchronize(s.ev_done));
	if (job.mode == MODE_FILL) {
		int64_t *ro = ctx->r_rec_off.as<int64_t>() + s.read0;
		for (int64_t r = 0; r < s.n_reads; ++r) ro[r] += s.bbase;
	} else {
		int64_t *uo = ctx->r_u_off.as<int64_t>() + s.read0, *bo = ctx->r_b_off.as<int64_t>() + s.read0;
		if (s.ubase || s.bbase)
			for (int64_t r = 0; r < s.n_reads; ++r) { uo[r] += s.ubase; bo[r] += s.bbase; }
	}
	s.state = 0;
	return MM2B200_OK;
}

// Encode one sub-batch into the slot's pinned staging buffer with the worker pool, upload it in one copy and restore
// a[] on the device.  *sent stays false when the anchors do not fit the wire format (the caller uploads them raw).
static int upload_wire(mm2b200_ctx *ctx, Slot &s, const mm128_t *src, int64_t na, bool *sent)
{
	int rc;
	const int64_t nblk = (na + WIRE_BLOCK - 1) >> WIRE_BLOCK_SHIFT;
	const size_t dict_cap = (size_t)na + 2 * WIRE_BLOCK;                       // beyond 4 B/anchor the format no longer pays
	const size_t pos_tab = (size_t)na * 8, pos_dict = pos_tab + (size_t)nblk * 8;
	if ((rc = s.h_a.reserve(pos_dict + dict_cap * 4))) return rc;
	char *h = s.h_a.as<char>();
	uint64_t *out = (uint64_t*)h, *tab = (uint64_t*)(h + pos_tab);
	uint32_t *dict = (uint32_t*)(h + pos_dict);
	int parts = (int)(nblk / 256);                                              // >= 64 Ki anchors per worker
	if (parts > ctx->pool.size()) parts = ctx->pool.size();
	if (parts < 1) parts = 1;
	if ((int)ctx->parts.size() < parts) ctx->parts.resize(parts);
	std::vector<WirePart> &pp = ctx->parts;
	const std::function<void(int)> work = [&](int k) {
		if (k >= parts) return;
		const int64_t b0 = nblk * k / parts, b1 = nblk * (k + 1) / parts;
		pp[k].ok = wire_encode(src, na, b0, b1, out, tab, pp[k].dict);
	};
	if (parts == 1) work(0); else ctx->pool.run(work);
	size_t total = 0;
	for (int k = 0; k < parts; ++k) {
		if (!pp[k].ok) return MM2B200_OK;
		total += pp[k].dict.size();
	}
	if (total > dict_cap) return MM2B200_OK;
	size_t base = 0;
	for (int k = 0; k < parts; ++k) {                                           // rebase the part-local dictionary positions
		const int64_t b0 = nblk * k / parts, b1 = nblk * (k + 1) / parts;
		if (base) for (int64_t b = b0; b < b1; ++b) tab[b] += base;
		if (!pp[k].dict.empty()) memcpy(dict + base, pp[k].dict.data(), pp[k].dict.size() * 4);
		base += pp[k].dict.size();
	}
	const size_t bytes = pos_dict + total * 4;
	// device staging: the compaction records (24 B/anchor) are dead until k_compact, long after the unpack
	CUDA_TRY(cudaMemcpyAsync((void*)s.w.rec, h, bytes, cudaMemcpyHostToDevice, s.stream));
	{ StageTimer t(ctx, s.stream, ST_PREP); t.done(launch_unpack_wire(s.w, (const void*)s.w.rec, pos_tab, pos_dict, s.stream)); }
	ctx->h2d_bytes += (int64_t)bytes;
	*sent = true;
	return MM2B200_OK;
}

static int run_batch(BatchJob &job)
{
	mm2b200_ctx *ctx = job.ctx;
	int rc;
	if ((rc = check_params(job.prm))) return rc;
	if (job.n_reads < 0 || (job.n_reads > 0 && (!job.offs || (!job.src && job.offs[job.n_reads] > job.offs[0]))))
		return fail(MM2B200_ERR_ARG, "bad batch arguments");
	for (int64_t r = 0; r < job.n_reads; ++r) {
		const int64_t n = job.offs[r + 1] - job.offs[r];
		if (n < 0) return fail(MM2B200_ERR_ARG, "offs[] is not non-decreasing at read %lld", (long long)r);
		if (n >= ((int64_t)1 << 29)) return fail(MM2B200_ERR_ARG, "read %lld has %lld anchors; the new_seed.p encoding (chain.c:95) needs n < 2^29", (long long)r, (long long)n);
	}
	std::lock_guard<std::mutex> lk(ctx->mu);
	CUDA_TRY(cudaSetDevice(ctx->device));
	const int64_t R = job.n_reads;
	ctx->h2d_bytes = 0; ctx->wire_sub_batches = 0;
	for (auto &s : ctx->slots)      // work that mm2b200_chain_device left running on a caller's stream still owns its slot
		if (s.user_pending) { CUDA_TRY(cudaEventSynchronize(s.ev_user)); s.user_pending = false; }
	// reserve the pinned result arenas to their upper bounds up front (sum n_v <= N, n_u <= n / max(min_cnt,1),
	// new_i <= n): a buffer must never move while an earlier chunk's D2H copy is still landing in it
	const int64_t N = R > 0 ? job.offs[R] - job.offs[0] : 0;
	if (job.mode == MODE_FILL) {
		if ((rc = ctx->r_rec_off.reserve((size_t)(R + 2) * 8))) return rc;
		if ((rc = ctx->r_rec.reserve((size_t)(N + 1) * 24))) return rc;
	} else {
		const int64_t mc = job.prm->min_cnt > 1 ? job.prm->min_cnt : 1;
		if ((rc = ctx->r_n_u.reserve((size_t)(R + 1) * 4))) return rc;
		if ((rc = ctx->r_u_off.reserve((size_t)(R + 2) * 8))) return rc;
		if ((rc = ctx->r_b_off.reserve((size_t)(R + 2) * 8))) return rc;
		if ((rc = ctx->r_u.reserve((size_t)(N / mc + R + 1) * 8))) return rc;
		if ((rc = ctx->r_b.reserve((size_t)(N + 1) * 16))) return rc;
	}
	job.src_pinned = job.src ? is_pinned(job.src) : true;
	// Long reads mean few runs per sub-batch (a 55 k-anchor read is one run): the fill kernel needs thousands of runs in
	// flight, so such batches keep more sub-batches on the device at once (their kernels overlap on separate streams).
	if (ctx->adaptive_slots && R > 0) {
		const int64_t avg = N / R > 0 ? N / R : 1;
		const int64_t reads_per_chunk = ctx->slot_anchors / avg > 0 ? ctx->slot_anchors / avg : 1;
		int64_t want = (4096 + reads_per_chunk - 1) / reads_per_chunk;
		const int64_t n_chunks_est = N / ctx->slot_anchors + 1;
		size_t free_b = 0, total_b = 0;
		cudaMemGetInfo(&free_b, &total_b);
		const int64_t mem_cap = (int64_t)(total_b / 2 / ((size_t)ctx->slot_anchors * 176 + 1));
		if (want > n_chunks_est) want = n_chunks_est;
		if (want > mem_cap) want = mem_cap;
		if (want > 32) want = 32;
		while ((int64_t)ctx->slots.size() < want) if ((rc = add_slot(ctx))) return rc;
	}
	const int S = (int)ctx->slots.size();
	int64_t r0 = 0, chunk = 0, next_a = 0;   // next_a: oldest chunk whose totals have not been collected yet
	while (r0 < R) {
		// greedy chunk: consecutive reads up to slot_anchors anchors (at least one read)
		int64_t r1 = r0 + 1;
		while (r1 < R && job.offs[r1 + 1] - job.offs[r0] <= ctx->slot_anchors && r1 - r0 < ((int64_t)1 << 24)) ++r1;
		const int64_t na = job.offs[r1] - job.offs[r0], nr = r1 - r0;
		Slot &s = ctx->slots[chunk % S];
		if (s.state == 1) {     // reuse: collect (in chunk order) everything up to the chunk that still owns this slot
			while (next_a <= chunk - S) { if ((rc = finish_a(job, ctx->slots[next_a % S]))) return rc; ++next_a; }
		}
		if (s.state == 2 && (rc = finish_b(job, s))) return rc;
		if ((rc = ensure_slot(ctx, s, (size_t)na, (size_t)nr))) return rc;
		if (job.mode == MODE_FILL && (rc = ensure_rec_out(s))) return rc;
		if ((rc = s.h_offs.reserve((size_t)(nr + 1) * 8))) return rc;
		if ((rc = s.h_totals.reserve(64))) return rc;
		int64_t *ho = s.h_offs.as<int64_t>();
		for (int64_t r = 0; r <= nr; ++r) ho[r] = job.offs[r0 + r] - job.offs[r0];
		s.read0 = r0; s.n_reads = nr; s.anchor0 = job.offs[r0] - job.offs[0]; s.n_anchors = na;
		Workspace &w = s.w;
		w.n_reads = (int32_t)nr; w.n_anchors = (int32_t)na;
		w.rec_out = (new_seed*)s.d_rec_out;
		CUDA_TRY(cudaMemcpyAsync((void*)w.offs, ho, (size_t)(nr + 1) * 8, cudaMemcpyHostToDevice, s.stream));
		void *dst = job.mode == MODE_BOTTOM ? (void*)w.rec : (void*)w.a;
		const char *src = job.src + (size_t)job.offs[r0] * job.elem;
		ctx->h2d_bytes += (nr + 1) * 8;
		if (na) {
			bool sent = false;
			if (ctx->wire && job.mode != MODE_BOTTOM) {
				if ((rc = upload_wire(ctx, s, (const mm128_t*)src, na, &sent))) return rc;
			}
			if (sent) ++ctx->wire_sub_batches;
			else {
				if (job.src_pinned) CUDA_TRY(cudaMemcpyAsync(dst, src, (size_t)na * job.elem, cudaMemcpyHostToDevice, s.stream));
				else {
					if ((rc = s.h_a.reserve((size_t)na * job.elem))) return rc;
					memcpy(s.h_a.ptr, src, (size_t)na * job.elem);
					CUDA_TRY(cudaMemcpyAsync(dst, s.h_a.ptr, (size_t)na * job.elem, cudaMemcpyHostToDevice, s.stream));
				}
				ctx->h2d_bytes += na * (int64_t)job.elem;
			}
		}
		if (job.mode == MODE_BOTTOM) {
			StageTimer t(ctx, s.stream, ST_COMPACT); t.done(launch_split_records(w, s.stream));
			enqueue_bottom(ctx, w, *job.prm, s.stream);
		} else {
			enqueue_fill(ctx, w, *job.prm, s.stream);
			enqueue_compact(ctx, w, *job.prm, s.stream);
			if (job.mode == MODE_CHAIN) enqueue_bottom(ctx, w, *job.prm, s.stream);
			else {
				StageTimer t(ctx, s.stream, ST_GATHER);
				int nl = launch_pack_records(w, s.stream);
				if (job.dbg_p || job.dbg_f || job.dbg_v) nl += launch_export_fp(w, s.stream);
				t.done(nl);
			}
		}
		CUDA_TRY(cudaMemcpyAsync(s.h_totals.ptr, w.totals, 64, cudaMemcpyDeviceToHost, s.stream));
		CUDA_TRY(cudaEventRecord(s.ev_totals, s.stream));
		CUDA_TRY(cudaGetLastError());
		s.state = 1;
		++chunk;
		r0 = r1;
		// move older chunks along without blocking: as soon as a chunk's totals are on the host, start its output copies.
		// With the default three slots the previous chunk is waited for (its D2H then overlaps this chunk's kernels).
		while (next_a < chunk - 1) {
			Slot &o = ctx->slots[next_a % S];
			if (S <= 3 || cudaEventQuery(o.ev_totals) == cudaSuccess) { if ((rc = finish_a(job, o))) return rc; ++next_a; }
			else break;
		}
	}
	// drain in chunk order
	while (next_a < chunk) { if ((rc = finish_a(job, ctx->slots[next_a % S]))) return rc; ++next_a; }
	for (int64_t c = chunk > S ? chunk - S : 0; c < chunk; ++c) {
		Slot &s = ctx->slots[c % S];
		if (s.state == 2 && (rc = finish_b(job, s))) return rc;
	}
	job.n_chunks = chunk;
	if (job.mode == MODE_FILL) ctx->r_rec_off.as<int64_t>()[R] = job.bbase;
	else { ctx->r_u_off.as<int64_t>()[R] = job.ubase; ctx->r_b_off.as<int64_t>()[R] = job.bbase; }
	return MM2B200_OK;
}

static void reset_slots(mm2b200_ctx *ctx)
{	// after an error: make sure no slot is left half-finished
	for (auto &s : ctx->slots) { if (s.stream) cudaStreamSynchronize(s.stream); s.state = 0; }
	cudaGetLastError();
}

extern "C" int mm2b200_chain_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *offs,
                                   const mm128_t *a, mm2b200_result_t *res)
{
	if (!ctx || !res) return fail(MM2B200_ERR_ARG, "NULL ctx or result");
	BatchJob job{};
	job.ctx = ctx; job.prm = prm; job.mode = MODE_CHAIN; job.n_reads = n_reads; job.offs = offs; job.src = (const char*)a; job.elem = 16;
	int rc = run_batch(job);
	if (rc) { reset_slots(ctx); return rc; }
	res->n_reads = n_reads;
	res->n_u = ctx->r_n_u.as<int32_t>(); res->u_off = ctx->r_u_off.as<int64_t>(); res->b_off = ctx->r_b_off.as<int64_t>();
	res->u = ctx->r_u.as<uint64_t>(); res->b = ctx->r_b.as<mm128_t>();
	res->n_units = job.n_units; res->n_sub_batches = job.n_chunks; res->n_visited = job.n_visited; res->n_lanes = job.n_lanes;
	return MM2B200_OK;
}

extern "C" int mm2b200_fill_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *offs,
                                  const mm128_t *a, mm2b200_records_t *out, int32_t *f, int32_t *p, int32_t *v)
{
	if (!ctx || !out) return fail(MM2B200_ERR_ARG, "NULL ctx or result");
	BatchJob job{};
	job.ctx = ctx; job.prm = prm; job.mode = MODE_FILL; job.n_reads = n_reads; job.offs = offs; job.src = (const char*)a; job.elem = 16;
	job.dbg_f = f; job.dbg_p = p; job.dbg_v = v;
	int rc = run_batch(job);
	if (rc) { reset_slots(ctx); return rc; }
	out->n_reads = n_reads;
	out->rec_off = ctx->r_rec_off.as<int64_t>();
	out->rec = ctx->r_rec.as<new_seed>();
	return MM2B200_OK;
}

extern "C" int mm2b200_bottom_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, const int64_t *rec_off,
                                    const struct new_seed *rec, mm2b200_result_t *res)
{
	if (!ctx || !res) return fail(MM2B200_ERR_ARG, "NULL ctx or result");
	BatchJob job{};
	job.ctx = ctx; job.prm = prm; job.mode = MODE_BOTTOM; job.n_reads = n_reads; job.offs = rec_off; job.src = (const char*)rec; job.elem = 24;
	int rc = run_batch(job);
	if (rc) { reset_slots(ctx); return rc; }
	res->n_reads = n_reads;
	res->n_u = ctx->r_n_u.as<int32_t>(); res->u_off = ctx->r_u_off.as<int64_t>(); res->b_off = ctx->r_b_off.as<int64_t>();
	res->u = ctx->r_u.as<uint64_t>(); res->b = ctx->r_b.as<mm128_t>();
	res->n_units = 0; res->n_sub_batches = job.n_chunks; res->n_visited = 0; res->n_lanes = 0;
	return MM2B200_OK;
}

extern "C" int mm2b200_partition_reads(int64_t n_reads, const int64_t *offs, int32_t n_parts, int64_t *cuts)
{
	if (n_reads < 0 || n_parts < 1 || !cuts || (n_reads > 0 && !offs)) return fail(MM2B200_ERR_ARG, "bad partition arguments");
	const int64_t base = n_reads ? offs[0] : 0, total = n_reads ? offs[n_reads] - base : 0;
	cuts[0] = 0;
	int64_t r = 0;
	for (int32_t k = 1; k < n_parts; ++k) {
		// first read whose start reaches k/n_parts of the anchors (ties resolved towards equal read counts when total == 0)
		const int64_t target = total ? (int64_t)(((__int128)total * k) / n_parts) : 0;
		if (total == 0) r = n_reads * k / n_parts;
		else while (r < n_reads && offs[r] - base < target) ++r;
		cuts[k] = r;
	}
	cuts[n_parts] = n_reads;
	return MM2B200_OK;
}

extern "C" int mm2b200_chain_batch_multi(mm2b200_ctx_t *const *ctxs, int32_t n_ctx, const mm2b200_params_t *prm, int64_t n_reads,
                                         const int64_t *offs, const mm128_t *a, int64_t *cuts, mm2b200_result_t *res)
{
	if (!ctxs || n_ctx < 1 || !cuts || !res) return fail(MM2B200_ERR_ARG, "bad multi-context arguments");
	int rc = mm2b200_partition_reads(n_reads, offs, n_ctx, cuts);
	if (rc) return rc;
	std::vector<int> rcs((size_t)n_ctx, 0);
	std::vector<std::string> errs((size_t)n_ctx);
	std::vector<std::thread> th;
	for (int32_t k = 0; k < n_ctx; ++k) {
		th.emplace_back([&, k] {
			const int64_t r0 = cuts[k], r1 = cuts[k + 1];
			// offs + r0 is a valid offsets array for the shard: the batch code rebases on its first entry
			rcs[k] = mm2b200_chain_batch(ctxs[k], prm, r1 - r0, offs + r0, a, &res[k]);
			if (rcs[k]) errs[k] = g_err;
		});
	}
	for (auto &t : th) t.join();
	for (int32_t k = 0; k < n_ctx; ++k)
		if (rcs[k]) { g_err = "shard " + std::to_string(k) + ": " + errs[k]; return rcs[k]; }
	return MM2B200_OK;
}

extern "C" int mm2b200_chain_device(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, int64_t n_anchors,
                                    const int64_t *d_offs, const mm128_t *d_a, void *stream, mm2b200_dev_result_t *out)
{
	int rc;
	if (!ctx || !out) return fail(MM2B200_ERR_ARG, "NULL ctx or result");
	if ((rc = check_params(prm))) return rc;
	if (n_reads < 0 || n_anchors < 0 || !d_offs || (n_anchors && !d_a)) return fail(MM2B200_ERR_ARG, "bad device batch arguments");
	std::lock_guard<std::mutex> lk(ctx->mu);
	CUDA_TRY(cudaSetDevice(ctx->device));
	Slot &s = ctx->slots[0];
	if (s.state) return fail(MM2B200_ERR_STATE, "slot 0 busy");
	cudaStream_t st = (cudaStream_t)stream;
	// back-to-back calls on one stream are ordered by the stream itself; a different stream must wait for the previous call
	if (s.user_pending && s.user_stream != st) { CUDA_TRY(cudaEventSynchronize(s.ev_user)); s.user_pending = false; }
	if ((rc = ensure_slot(ctx, s, (size_t)n_anchors, (size_t)n_reads))) return rc;
	Workspace w = s.w;              // private copy pointing at the caller's inputs
	w.a = (const ulonglong2*)d_a; w.offs = d_offs;
	w.n_reads = (int32_t)n_reads; w.n_anchors = (int32_t)n_anchors;
	if (ctx->split_device && n_reads >= 4096) {
		// Two halves of the reads on two streams: the HBM-bound stages of one half (prep, compaction, chain ends / backtrack)
		// run beside the issue-bound fill of the other.  Per-anchor arrays are shared (anchor indices are batch-wide), per-read
		// arrays are offset views, each half has its own run list and counters; offsets, totals and the gather run once
		// over all reads after the join.
		const int64_t r1 = n_reads / 2;
		Workspace h1 = w, h2 = w;
		h1.half = 1; h1.n_reads = (int32_t)r1;
		h2.half = 2; h2.n_reads = (int32_t)(n_reads - r1);
		h2.offs = w.offs + r1; h2.avg = w.avg + r1; h2.new_i = w.new_i + r1; h2.n_u = w.n_u + r1; h2.n_v = w.n_v + r1;
		h2.units = w.units + w.units_cap; h2.counters = w.counters + CNT_N;
		CUDA_TRY(cudaMemsetAsync(w.visited, 0, 2 * sizeof(unsigned long long), st));
		CUDA_TRY(cudaEventRecord(ctx->ev_fork, st));
		CUDA_TRY(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0));
		const Workspace *hv[2] = { &h1, &h2 };
		cudaStream_t sv[2] = { st, ctx->aux_stream };
		for (int k = 0; k < 2; ++k) {
			enqueue_fill(ctx, *hv[k], *prm, sv[k]);
			enqueue_compact(ctx, *hv[k], *prm, sv[k]);
			{ StageTimer t(ctx, sv[k], ST_TAIL); t.done(launch_tail_reads(*hv[k], *prm, sv[k])); }
		}
		CUDA_TRY(cudaEventRecord(ctx->ev_join, ctx->aux_stream));
		CUDA_TRY(cudaStreamWaitEvent(st, ctx->ev_join, 0));
		{ StageTimer t(ctx, st, ST_TAIL); t.done(launch_tail_scan(w, st)); }
		{ StageTimer t(ctx, st, ST_GATHER); t.done(launch_gather(w, st)); }
	} else {
		enqueue_fill(ctx, w, *prm, st);
		enqueue_compact(ctx, w, *prm, st);
		enqueue_bottom(ctx, w, *prm, st);
	}
	CUDA_TRY(cudaGetLastError());
	CUDA_TRY(cudaEventRecord(s.ev_user, st));
	s.user_stream = st; s.user_pending = true;
	out->d_n_u = w.n_u; out->d_u_off = w.u_off; out->d_b_off = w.b_off;
	out->d_u = w.u_out; out->d_b = (const mm128_t*)w.b_out; out->d_totals = w.totals;
	return MM2B200_OK;
}

// what seed_api.cu needs from a context without seeing its layout
namespace mm2b200 {
int api_fail(int code, const char *msg) { return fail(code, "%s", msg); }
int ctx_device(const mm2b200_ctx *ctx) { return ctx->device; }
int ctx_profile(const mm2b200_ctx *ctx) { return ctx->profile; }
int ctx_sm_count(const mm2b200_ctx *ctx) { return ctx->lc.sm_count; }
void *ctx_ext(const mm2b200_ctx *ctx, int k) { return ctx->ext[k]; }
void ctx_set_ext(mm2b200_ctx *ctx, int k, void *p, void (*free_fn)(void*)) { ctx->ext[k] = p; ctx->ext_free[k] = free_fn; }
void ctx_add_launches(mm2b200_ctx *ctx, int64_t n) { ctx->launches.fetch_add(n); }
}

extern "C" int mm2b200_fill_device(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, int64_t n_reads, int64_t n_anchors,
                                   const int64_t *d_offs, const mm128_t *d_a, void *stream, mm2b200_dev_records_t *out)
{
	int rc;
	if (!ctx || !out) return fail(MM2B200_ERR_ARG, "NULL ctx or result");
	if ((rc = check_params(prm))) return rc;
	if (n_reads < 0 || n_anchors < 0 || !d_offs || (n_anchors && !d_a)) return fail(MM2B200_ERR_ARG, "bad device batch arguments");
	std::lock_guard<std::mutex> lk(ctx->mu);
	CUDA_TRY(cudaSetDevice(ctx->device));
	Slot &s = ctx->slots[0];
	if (s.state) return fail(MM2B200_ERR_STATE, "slot 0 busy");
	cudaStream_t st = (cudaStream_t)stream;
	if (s.user_pending && s.user_stream != st) { CUDA_TRY(cudaEventSynchronize(s.ev_user)); s.user_pending = false; }
	if ((rc = ensure_slot(ctx, s, (size_t)n_anchors, (size_t)n_reads))) return rc;
	if (s.cap_rec_out < s.cap_anchors || !s.d_rec_out) {
		// the record buffer may still be read by the previous call on this stream
		if (s.user_pending) { CUDA_TRY(cudaEventSynchronize(s.ev_user)); s.user_pending = false; }
		if ((rc = ensure_rec_out(s))) return rc;
	}
	Workspace w = s.w;
	w.a = (const ulonglong2*)d_a; w.offs = d_offs;
	w.n_reads = (int32_t)n_reads; w.n_anchors = (int32_t)n_anchors;
	w.rec_out = (new_seed*)s.d_rec_out;
	enqueue_fill(ctx, w, *prm, st);
	enqueue_compact(ctx, w, *prm, st);
	{ StageTimer t(ctx, st, ST_GATHER); t.done(launch_pack_records(w, st)); }
	CUDA_TRY(cudaGetLastError());
	CUDA_TRY(cudaEventRecord(s.ev_user, st));
	s.user_stream = st; s.user_pending = true;
	out->d_rec_off = w.rec_off; out->d_rec = w.rec_out; out->d_totals = w.totals;
	return MM2B200_OK;
}

extern "C" int mm2b200_int_peak(mm2b200_ctx_t *ctx, int mode, double *lane_ops_per_s)
{
	if (!ctx || !lane_ops_per_s) return fail(MM2B200_ERR_ARG, "NULL argument");
	std::lock_guard<std::mutex> lk(ctx->mu);
	CUDA_TRY(cudaSetDevice(ctx->device));
	unsigned *d_out = nullptr;
	CUDA_TRY(cudaMalloc(&d_out, 64));
	cudaEvent_t e0, e1;
	CUDA_TRY(cudaEventCreate(&e0)); CUDA_TRY(cudaEventCreate(&e1));
	cudaStream_t st = ctx->slots[0].stream;
	const int grid = ctx->lc.sm_count * 8;      // 8 CTAs x 256 threads = 64 warps per SM
	launch_int_peak(mode, grid, 256, d_out, st);   // warm-up
	double best = 0;
	for (int rep = 0; rep < 5; ++rep) {
		cudaEventRecord(e0, st);
		const long long ops = launch_int_peak(mode, grid, 4096, d_out, st);
		cudaEventRecord(e1, st);
		CUDA_TRY(cudaEventSynchronize(e1));
		float ms = 0;
		cudaEventElapsedTime(&ms, e0, e1);
		ctx->launches += 1;
		if (ms > 0 && ops / (ms * 1e-3) > best) best = ops / (ms * 1e-3);
	}
	cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d_out);
	CUDA_TRY(cudaGetLastError());
	*lane_ops_per_s = best;
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Section 1: reference entry points.  Concurrent callers of mm_chain_dp (kt_for workers, map.c:598-604), of
// mm_chain_dp_fpga (map.c:525,941) and of mm_chain_dp_bottom (map.c:837) are combined by one dispatcher thread into
// device batches, one per (entry point, parameter set): while the GPU works on one batch the next callers queue up,
// so the batch size adapts to the number of host threads without any timer.  The dispatcher owns a private context:
// nothing else ever touches its result arenas, and every caller's slice is copied out before the next batch starts.
//
// Errors.  The reference reports none and "NULL, n_u = 0" is also the legitimate "no chain" answer, so a
// failure must not look like an unmapped read: it is printed to stderr, and unless MM2B200_ON_ERROR=null the process
// aborts (default; wrong PAF is never produced silently).  With MM2B200_ON_ERROR=null the call returns NULL / 0 and
// mm2b200_last_error() tells the two apart.  A batch that fails is retried one call at a time, so one bad read (or a
// transient allocation failure on a big batch) cannot fail its neighbours.
// ---------------------------------------------------------------------------------------------------------
namespace {

enum EntryMode { ENTRY_CHAIN = 0, ENTRY_FPGA = 1, ENTRY_BOTTOM = 2 };

struct Pending {
	int mode = ENTRY_CHAIN;
	mm2b200_params_t prm;
	int64_t n = 0;
	const void *in = nullptr;     // mm128_t[n] (ENTRY_CHAIN, ENTRY_FPGA) or new_seed[n] (ENTRY_BOTTOM)
	// filled by the dispatcher: malloc()ed copies of this call's slice of the batch result, so the dispatcher never has
	// to wait for the callers before it starts the next batch
	int rc = 0; std::string err;
	int32_t n_u = 0; uint64_t *u = nullptr; mm128_t *b = nullptr; int64_t n_b = 0;
	struct new_seed *rec = nullptr; int64_t n_rec = 0;
	bool done = false;
	std::condition_variable cv;   // one per call: only the owner is woken
};

struct Batcher {
	std::mutex mu;
	std::condition_variable cv_work;
	std::deque<Pending*> q;
	// Two dispatchers, each with a private context: while one batch is on the device the other dispatcher assembles and
	// launches the next (the fork's worker threads sit in mm_chain_dp_fpga while its result threads sit in mm_chain_dp_bottom).
	static const int N_DISPATCH = 2;
	mm2b200_ctx *ctx[N_DISPATCH] = {nullptr, nullptr};
	std::thread th[N_DISPATCH];
	int init_rc = 0; std::string init_err;
	bool started = false, stop = false;
	std::atomic<int64_t> n_batches{0}, n_calls{0};
};

Batcher *g_batcher = nullptr;
std::once_flag g_batcher_once;

bool abort_on_error()
{
	static const bool v = [] { const char *e = getenv("MM2B200_ON_ERROR"); return !(e && !strcmp(e, "null")); }();
	return v;
}

// a drop-in symbol failed: never silently (see the section comment)
void entry_failed(const char *sym, const std::string &msg)
{
	g_err = msg;
	fprintf(stderr, "[mm2chain-b200] %s failed: %s%s\n", sym, msg.c_str(),
	        abort_on_error() ? " -- aborting (MM2B200_ON_ERROR=null returns NULL and keeps mm2b200_last_error() instead)" : "");
	if (abort_on_error()) abort();
}

// one device batch over calls[0..m): all share mode and parameters.  Returns the batch's rc; on success every call has
// its private result copy.
int run_entry_batch(mm2b200_ctx *ctx, PinnedBuf &stage, std::vector<int64_t> &offs, Pending *const *calls, size_t m)
{
	const int mode = calls[0]->mode;
	const size_t elem = mode == ENTRY_BOTTOM ? sizeof(struct new_seed) : sizeof(mm128_t);
	int64_t total = 0;
	offs.assign(1, 0);
	for (size_t k = 0; k < m; ++k) { total += calls[k]->n; offs.push_back(total); }
	int rc = stage.reserve((size_t)(total > 0 ? total : 1) * elem);
	if (rc) return rc;
	char *dst = stage.as<char>();
	for (size_t k = 0; k < m; ++k)
		if (calls[k]->n > 0) memcpy(dst + (size_t)offs[k] * elem, calls[k]->in, (size_t)calls[k]->n * elem);
	if (mode == ENTRY_FPGA) {
		mm2b200_records_t out{};
		if ((rc = mm2b200_fill_batch(ctx, &calls[0]->prm, (int64_t)m, offs.data(), (const mm128_t*)dst, &out, nullptr, nullptr, nullptr))) return rc;
		for (size_t k = 0; k < m; ++k) {
			Pending *p = calls[k];
			p->n_rec = out.rec_off[k + 1] - out.rec_off[k];
			p->rec = (struct new_seed*)malloc((size_t)(p->n_rec > 0 ? p->n_rec : 1) * sizeof(struct new_seed));   // chain.c:233: libc malloc
			if (!p->rec) return fail(MM2B200_ERR_NOMEM, "malloc of %lld records failed", (long long)p->n_rec);
			if (p->n_rec > 0) memcpy(p->rec, out.rec + out.rec_off[k], (size_t)p->n_rec * sizeof(struct new_seed));
		}
		return MM2B200_OK;
	}
	mm2b200_result_t res{};
	if (mode == ENTRY_BOTTOM) rc = mm2b200_bottom_batch(ctx, &calls[0]->prm, (int64_t)m, offs.data(), (const struct new_seed*)dst, &res);
	else rc = mm2b200_chain_batch(ctx, &calls[0]->prm, (int64_t)m, offs.data(), (const mm128_t*)dst, &res);
	if (rc) return rc;
	for (size_t k = 0; k < m; ++k) {
		Pending *p = calls[k];
		if (res.n_u[k] <= 0) continue;
		p->n_u = res.n_u[k];
		p->n_b = res.b_off[k + 1] - res.b_off[k];
		p->u = (uint64_t*)malloc((size_t)p->n_u * 8);
		p->b = (mm128_t*)malloc((size_t)(p->n_b > 0 ? p->n_b : 1) * sizeof(mm128_t));
		if (!p->u || !p->b) return fail(MM2B200_ERR_NOMEM, "malloc of a chain result (%d chains, %lld anchors) failed", p->n_u, (long long)p->n_b);
		memcpy(p->u, res.u + res.u_off[k], (size_t)p->n_u * 8);
		memcpy(p->b, res.b + res.b_off[k], (size_t)p->n_b * sizeof(mm128_t));
	}
	return MM2B200_OK;
}

void drop_results(Pending *p)
{
	free(p->u); free(p->b); free(p->rec);
	p->u = nullptr; p->b = nullptr; p->rec = nullptr; p->n_u = 0; p->n_b = 0; p->n_rec = 0;
}

void dispatcher_main(Batcher *B, int which)
{
	mm2b200_ctx *ctx = B->ctx[which];
	std::vector<int64_t> offs;
	PinnedBuf stage;
	size_t recent = 0;                  // calls in this dispatcher's previous batch
	for (;;) {
		std::vector<Pending*> batch;
		{
			std::unique_lock<std::mutex> lk(B->mu);
			B->cv_work.wait(lk, [&] { return B->stop || !B->q.empty(); });
			if (B->q.empty()) break;     // stop requested and nothing left to serve
			// The callers of the previous batch are on their way back (they were woken a few microseconds ago): when only a
			// fraction of them has queued again, wait briefly instead of spending a device round trip on a handful of reads.
			if (B->q.size() * 2 < recent)
				B->cv_work.wait_for(lk, std::chrono::microseconds(150), [&] { return B->stop || B->q.size() * 2 >= recent; });
			if (B->q.empty()) continue;  // the other dispatcher took them
			// the entry point with the most queued calls; of those, every call that shares the first one's parameters
			size_t cnt[3] = {0, 0, 0};
			for (auto *p : B->q) ++cnt[p->mode];
			const int mode = cnt[1] > cnt[0] ? (cnt[2] > cnt[1] ? 2 : 1) : (cnt[2] > cnt[0] ? 2 : 0);
			const Pending *key = nullptr;
			for (auto it = B->q.begin(); it != B->q.end();) {
				if ((*it)->mode == mode && (!key || !memcmp(&(*it)->prm, &key->prm, sizeof(key->prm)))) {
					if (!key) key = *it;
					batch.push_back(*it); it = B->q.erase(it);
				} else ++it;
			}
			if (!B->q.empty()) B->cv_work.notify_one();   // something else is left for the other dispatcher
		}
		recent = batch.size();
		++B->n_batches; B->n_calls += (int64_t)batch.size();
		int rc = run_entry_batch(ctx, stage, offs, batch.data(), batch.size());
		if (rc && batch.size() > 1) {
			// retry one call at a time: a single bad read must not fail the reads that happened to share its batch
			for (auto *p : batch) {
				drop_results(p);
				p->rc = run_entry_batch(ctx, stage, offs, &p, 1);
				if (p->rc) { p->err = g_err; drop_results(p); }
			}
		} else for (auto *p : batch) { p->rc = rc; if (rc) { p->err = g_err; drop_results(p); } }
		{
			std::lock_guard<std::mutex> lk(B->mu);
			for (auto *p : batch) { p->done = true; p->cv.notify_one(); }
		}
	}
	stage.release();
}

Batcher *batcher()
{
	std::call_once(g_batcher_once, [] {
		Batcher *B = new Batcher();
		mm2b200_config_t cfg{};
		const char *dev = getenv("MM2B200_DEVICE");
		cfg.device = dev ? atoi(dev) : 0;
		cfg.n_slots = 2;
		cfg.slot_anchors = (int64_t)4 << 20;
		for (int k = 0; k < Batcher::N_DISPATCH && !B->init_rc; ++k)
			if (mm2b200_create(&cfg, &B->ctx[k]) != MM2B200_OK) { B->init_rc = MM2B200_ERR_CUDA; B->init_err = g_err.empty() ? "mm2b200_create failed" : g_err; }
		if (!B->init_rc) {
			for (int k = 0; k < Batcher::N_DISPATCH; ++k) B->th[k] = std::thread(dispatcher_main, B, k);
			B->started = true;
		}
		g_batcher = B;
	});
	return g_batcher;
}

// library unload / process exit: let the dispatcher finish what is queued, then join it.  (The context itself is left to
// the CUDA runtime's own teardown: calling into CUDA from a destructor at exit is not safe.)
__attribute__((destructor)) void mm2b200_on_unload()
{
	Batcher *B = g_batcher;
	if (!B || !B->started) return;
	{ std::lock_guard<std::mutex> lk(B->mu); B->stop = true; }
	B->cv_work.notify_all();
	for (auto &t : B->th) if (t.joinable()) t.join();
	B->started = false;
}

// submit one call and wait for it; returns false (after reporting) when the call failed
bool submit_entry(const char *sym, Pending &pd)
{
	Batcher *B = batcher();
	if (!B->started) { entry_failed(sym, B->init_err.empty() ? "the dispatcher is not running" : B->init_err); return false; }
	{
		std::unique_lock<std::mutex> lk(B->mu);
		B->q.push_back(&pd);
		B->cv_work.notify_one();
		pd.cv.wait(lk, [&] { return pd.done; });
	}
	if (pd.rc != MM2B200_OK) { entry_failed(sym, pd.err); return false; }
	return true;
}

void *host_alloc(void *km, size_t bytes)
{
	if (kmalloc) return kmalloc(km, bytes);
	return malloc(bytes ? bytes : 1);
}
void host_free(void *km, void *p)
{
	if (kfree) kfree(km, p);
	else free(p);
}

} // namespace

extern "C" int mm2b200_entry_stats(int64_t *n_batches, int64_t *n_calls)
{
	Batcher *B = g_batcher;
	if (n_batches) *n_batches = B ? B->n_batches.load() : 0;
	if (n_calls) *n_calls = B ? B->n_calls.load() : 0;
	return MM2B200_OK;
}

extern "C" mm128_t *mm_chain_dp(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_cnt, int min_sc, int is_cdna,
                                int n_segs, int64_t n, mm128_t *a, int *n_u_, uint64_t **_u, void *km)
{
	if (_u) *_u = 0;
	if (n_u_) *n_u_ = 0;
	g_err.clear();
	if (km && !kmalloc) { entry_failed("mm_chain_dp", "km != NULL but the host program exports no kmalloc/kfree"); return nullptr; }
	mm128_t *ret = nullptr;
	if (n > 0) {
		Pending pd;
		pd.mode = ENTRY_CHAIN;
		pd.prm = mm2b200_params_t{max_dist_x, max_dist_y, bw, max_skip, min_cnt, min_sc, is_cdna, n_segs, 0};
		pd.n = n; pd.in = a;
		if (submit_entry("mm_chain_dp", pd) && pd.n_u > 0) {
			uint64_t *u = pd.u;
			ret = pd.b;
			if (km) {            // results must live in the caller's kalloc arena: copy on the caller's thread
				// (km == NULL: kmalloc(NULL, n) is malloc(n) (kalloc.c:127), so the malloc()ed copies are handed over as they are)
				u = (uint64_t*)host_alloc(km, (size_t)pd.n_u * 8);
				ret = (mm128_t*)host_alloc(km, (size_t)pd.n_b * sizeof(mm128_t));
				memcpy(u, pd.u, (size_t)pd.n_u * 8);
				memcpy(ret, pd.b, (size_t)pd.n_b * sizeof(mm128_t));
				free(pd.u); free(pd.b);
			}
			if (_u) *_u = u; else host_free(km, u);
			if (n_u_) *n_u_ = pd.n_u;
		}
	}
	host_free(km, a);   // chain.c:137,214: the input array is consumed
	return ret;
}

extern "C" struct new_seed *mm_chain_dp_fpga(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_sc, int is_cdna,
                                             int n_segs, int64_t n, mm128_t *a, uint32_t *_new_i)
{
	if (_new_i) *_new_i = 0;
	g_err.clear();
	struct new_seed *ret = nullptr;
	if (n > 0) {
		Pending pd;
		pd.mode = ENTRY_FPGA;
		pd.prm = mm2b200_params_t{max_dist_x, max_dist_y, bw, max_skip, 0, min_sc, is_cdna, n_segs, 0};
		pd.n = n; pd.in = a;
		if (submit_entry("mm_chain_dp_fpga", pd)) {
			ret = pd.rec;
			if (_new_i) *_new_i = (uint32_t)pd.n_rec;
		}
	} else ret = (struct new_seed*)malloc(sizeof(struct new_seed));   // chain.c:233 with n == 0
	free(a);   // chain.c:322
	return ret;
}

extern "C" mm128_t *mm_chain_dp_bottom(int min_cnt, int min_sc, int n_segs, int *n_u_, uint64_t **_u, void *km,
                                       struct new_seed *fpga_a, uint32_t new_i)
{
	(void)km;   // the reference ignores km here as well (chain.c:341-358,400: malloc / free)
	if (_u) *_u = 0;
	if (n_u_) *n_u_ = 0;
	g_err.clear();
	if (new_i == 0) return nullptr;
	Pending pd;
	pd.mode = ENTRY_BOTTOM;
	pd.prm = mm2b200_params_t{0, 0, 0, 0, min_cnt, min_sc, 0, n_segs, 0};
	pd.n = (int64_t)new_i; pd.in = fpga_a;
	if (!submit_entry("mm_chain_dp_bottom", pd) || pd.n_u <= 0) return nullptr;
	if (_u) *_u = pd.u; else free(pd.u);
	if (n_u_) *n_u_ = pd.n_u;
	return pd.b;
}
