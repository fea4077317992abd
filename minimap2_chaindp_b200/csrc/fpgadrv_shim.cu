// fpgadrv_shim.cu -- the fork's driver boundary (fpga.h:37-62) on B200s: the 11 symbols the
// fork's objects expect from the closed libfpgadrv.a, implemented over the seeding + chaining device path of this library.
// See include/mm2fpgadrv_b200.h for the contract.  Host code only: the device work is in seed_api.cu (seed_packets_begin /
// seed_packets_end; mm2b200_seed_fill_batch on the careful path).
//
// Structure (B200-first, not the FPGA's ring of fixed buffers):
//   fpga_get_writebuf     -> a page-locked, recycled buffer the fork fills with one request packet
//   fpga_writebuf_submit  -> a queue of request packets
//   workers (2 per GPU)   -> take everything that waits (bounded) as ONE device batch.  Direct path: every packet is DMA'd
//                            as it is, unpacked on the device, and the response packets are laid out in HBM and DMA'd into
//                            page-locked response buffers; the host only parses headers and computes the response layout.
//                            Careful path (a batch the device refused as a whole, or MM2B200_SHIM_ZEROCOPY=0): minimizers
//                            copied into pinned staging, records / mini_pos cut back into response packets by the worker.
//   fpga_get_retbuf       -> a queue of finished response packets
// On the careful path a refused batch (allocation failure, anchor count beyond the index space) is halved and retried; a
// single read that still fails is answered with err_flag = 1, the fork's own "host, redo this read" protocol (map.c:933-944).
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#include "mm2fpgadrv_b200.h"
#include "seed_internal.h"
#include "packet_format.h"

using namespace mm2b200;

namespace {

struct Packet { void *buf; uint32_t size; };

struct ShimDevice {
	int ordinal = 0;
	mm2b200_ctx_t *ctx = nullptr;
	mm2b200_index_t *idx = nullptr;
	bool index_dirty = false;          // streams loaded since the last commit (guarded by Shim::mu)
	std::thread th;
	// pinned staging of one batch (owned by the worker)
	mm128_t *mv = nullptr; size_t mv_cap = 0;
};

struct Shim {
	std::mutex mu;
	std::condition_variable cv_work, cv_ret, cv_idle;
	std::deque<Packet> pending;
	std::deque<Packet> done;
	int64_t pending_bytes = 0;         // request bytes handed out by get_writebuf and not yet worked off
	int busy = 0;                      // workers inside a batch
	bool exit_block = false, stop = false;
	bool have_params = false, have_index = false;
	bool index_open = false;           // stream chunks are arriving; closed by fpga_set_params or the first packet
	mm2b200_seed_params_t prm{};
	std::vector<std::unique_ptr<ShimDevice>> dev;
	int64_t max_bytes = (int64_t)1 << 30;
	int64_t batch_mv = (int64_t)8 << 20;
	bool direct = true;                // packets are DMA'd as they are (MM2B200_SHIM_ZEROCOPY=0: staged and cut by the worker)
	std::atomic<int64_t> n_packets{0}, n_reads{0}, n_batches{0}, n_err{0};
};

// Packet buffers (requests and responses, a few hundred KB each) come from size-classed free lists: a fresh malloc of that
// size is an mmap whose pages fault in on first touch -- measured, that cost more than the device batch itself.  The buffers
// are PAGE-LOCKED (cudaHostAlloc, portable): the fork fills a request packet in place and the GPU's copy engine reads it
// from there; response packets are written by the copy engine and handed to the fork as they are.  Ordinary memory is the
// fallback when page-locked memory runs out (the copies then go through the driver's own staging) or with
// MM2B200_SHIM_ZEROCOPY=0.
struct BufPool {
	static const int N_CLASS = 24;                 // 4 KB << c
	std::mutex mu;
	std::vector<void*> free_list[N_CLASS];
	int64_t pooled = 0, cap = (int64_t)2 << 30;
	int device = 0;
	bool pin = true;
	// 16-byte prefix: [0] size class (-1: larger than the largest class), [1] 1 = page-locked
	static int cls(size_t n) { int c = 0; while (((size_t)4096 << c) < n + 16 && c < N_CLASS - 1) ++c; return c; }
	char *raw_alloc(size_t bytes, int64_t c)
	{
		char *q = nullptr;
		int64_t pinned = 0;
		if (pin) {
			int cur = -1;
			if (cudaGetDevice(&cur) == cudaSuccess && cur != device) cudaSetDevice(device);
			if (cudaHostAlloc((void**)&q, bytes, cudaHostAllocPortable) == cudaSuccess) pinned = 1;
			else { cudaGetLastError(); q = nullptr; }
		}
		if (!q) q = (char*)malloc(bytes);
		if (!q) return nullptr;
		((int64_t*)q)[0] = c; ((int64_t*)q)[1] = pinned;
		return q;
	}
	static void raw_free(char *q) { if (((int64_t*)q)[1]) cudaFreeHost(q); else free(q); }
	void *get(size_t n)
	{
		const int c = cls(n);
		if (((size_t)4096 << c) < n + 16) {            // larger than the largest class: not recycled
			char *q = raw_alloc(n + 16, -1);
			return q ? q + 16 : nullptr;
		}
		{
			std::lock_guard<std::mutex> lk(mu);
			if (!free_list[c].empty()) { void *q = free_list[c].back(); free_list[c].pop_back(); pooled -= (int64_t)4096 << c; return (char*)q + 16; }
		}
		char *q = raw_alloc((size_t)4096 << c, c);
		return q ? q + 16 : nullptr;
	}
	void put(void *user)
	{
		if (!user) return;
		char *q = (char*)user - 16;
		const int64_t c = *(int64_t*)q;
		if (c >= 0) {
			std::lock_guard<std::mutex> lk(mu);
			if (pooled + ((int64_t)4096 << c) <= cap) { free_list[c].push_back(q); pooled += (int64_t)4096 << c; return; }
		}
		raw_free(q);
	}
	void drain()
	{
		std::lock_guard<std::mutex> lk(mu);
		for (auto &l : free_list) { for (void *q : l) raw_free((char*)q); l.clear(); }
		pooled = 0;
	}
};
BufPool g_pool;

Shim *g_shim = nullptr;
std::mutex g_shim_mu;                  // guards creation / destruction of g_shim

bool shim_abort_on_error()
{
	const char *e = getenv("MM2B200_ON_ERROR");
	return !(e && !strcmp(e, "null"));
}

void shim_fatal(const char *what, const char *msg)
{
	fprintf(stderr, "[mm2chain-b200] driver shim: %s: %s -- aborting\n", what, msg ? msg : "");
	abort();
}

int64_t env_i64(const char *name, int64_t dflt)
{
	const char *e = getenv(name);
	if (!e || !*e) return dflt;
	char *end = nullptr;
	long long v = strtoll(e, &end, 10);
	if (end && (*end == 'k' || *end == 'K')) v <<= 10;
	else if (end && (*end == 'm' || *end == 'M')) v <<= 20;
	else if (end && (*end == 'g' || *end == 'G')) v <<= 30;
	return v > 0 ? (int64_t)v : dflt;
}

std::vector<int> shim_devices()
{
	std::vector<int> out;
	const int n = mm2b200_device_count();
	const char *e = getenv("MM2B200_DEVICES");
	if (e && !strcmp(e, "all")) {
		for (int i = 0; i < n; ++i) out.push_back(i);
	} else if (e && *e) {
		const char *p = e;
		while (*p) {
			char *end = nullptr;
			long v = strtol(p, &end, 10);
			if (end == p) break;
			if (v >= 0 && v < n) out.push_back((int)v);
			p = *end == ',' ? end + 1 : end;
		}
	}
	if (out.empty()) {
		const char *d = getenv("MM2B200_DEVICE");
		out.push_back(d ? atoi(d) : 0);
	}
	return out;
}

// ---- one read of a batch ----------------------------------------------------------------------------------------
struct ReadRef {
	int pkt, slot;                     // request packet and position inside it
	PktTask task;
	const char *mv;                    // minimizers inside the request packet
	// result: either a view into the index object's arenas (valid until its next call) or owned copies
	const struct new_seed *rec = nullptr; int64_t n_rec = 0;
	const uint64_t *mini = nullptr; int64_t n_mini = 0;
	int32_t rep_len = 0;
	bool err = false, in_arena = false;
	std::vector<char> keep;
};

// results of the previous device call still live in the arenas: copy them out before the next call overwrites them
void evacuate(std::vector<ReadRef> &reads, const std::vector<int> &members)
{
	for (int k : members) {
		ReadRef &r = reads[(size_t)k];
		if (!r.in_arena) continue;
		const size_t rb = (size_t)r.n_rec * sizeof(struct new_seed), mb = (size_t)r.n_mini * 8;
		r.keep.resize(rb + mb + 8);
		if (rb) memcpy(r.keep.data(), r.rec, rb);
		char *m = r.keep.data() + ((rb + 7) & ~(size_t)7);
		if (mb) memcpy(m, r.mini, mb);
		r.rec = (const struct new_seed*)r.keep.data();
		r.mini = (const uint64_t*)m;
		r.in_arena = false;
	}
}

bool ensure_stage(ShimDevice &d, size_t n_mv)
{
	if (n_mv <= d.mv_cap && d.mv) return true;
	if (d.mv) mm2b200_pinned_free(d.mv);
	d.mv_cap = n_mv + n_mv / 4 + 1024;
	d.mv = (mm128_t*)mm2b200_pinned_alloc(d.mv_cap * sizeof(mm128_t));
	if (!d.mv) { d.mv_cap = 0; return false; }
	return true;
}

// one device call over reads members[lo..hi) (all share gap_ref / gap_qry); false = the device refused
bool run_range(Shim &S, ShimDevice &d, const mm2b200_seed_params_t &prm, std::vector<ReadRef> &reads, const std::vector<int> &members,
               size_t lo, size_t hi, std::vector<int> &last_call, std::string &err)
{
	evacuate(reads, last_call);
	last_call.clear();
	size_t n_mv = 0;
	for (size_t k = lo; k < hi; ++k) n_mv += (size_t)reads[(size_t)members[k]].task.seednum;
	if (!ensure_stage(d, n_mv + 1)) { err = "pinned staging allocation failed"; return false; }
	const size_t R = hi - lo;
	std::vector<int64_t> off(R + 1, 0);
	std::vector<uint32_t> bid(R);
	std::vector<int32_t> qlen(R);
	size_t at = 0;
	for (size_t k = lo; k < hi; ++k) {
		const ReadRef &r = reads[(size_t)members[k]];
		const size_t n = (size_t)r.task.seednum;
		if (n) memcpy(d.mv + at, r.mv, n * sizeof(mm128_t));
		at += n;
		off[k - lo + 1] = (int64_t)at;
		bid[k - lo] = r.task.bid;
		qlen[k - lo] = r.task.qlensum;
	}
	mm2b200_seed_records_t rr;
	memset(&rr, 0, sizeof(rr));
	const int rc = mm2b200_seed_fill_batch(d.idx, &prm, (int64_t)R, off.data(), d.mv, bid.data(), qlen.data(), &rr);
	if (rc != MM2B200_OK) {
		err = mm2b200_last_error();
		// a CUDA / state error is not a property of the reads: never turn it into silent host work unless asked to
		if ((rc == MM2B200_ERR_CUDA || rc == MM2B200_ERR_STATE) && shim_abort_on_error()) shim_fatal("device batch failed", err.c_str());
		return false;
	}
	S.n_batches.fetch_add(1);
	for (size_t k = lo; k < hi; ++k) {
		ReadRef &r = reads[(size_t)members[k]];
		const size_t q = k - lo;
		r.rec = rr.rec + rr.rec_off[q]; r.n_rec = rr.rec_off[q + 1] - rr.rec_off[q];
		r.mini = rr.mini_pos + rr.mini_off[q]; r.n_mini = rr.mini_off[q + 1] - rr.mini_off[q];
		r.rep_len = rr.rep_len[q];
		r.in_arena = true; r.err = false;
		last_call.push_back(members[k]);
	}
	return true;
}

// all reads of one (gap_ref, gap_qry) group: whole group first, halves on refusal, err_flag for a single refused read
void run_group(Shim &S, ShimDevice &d, const mm2b200_seed_params_t &prm, std::vector<ReadRef> &reads, const std::vector<int> &members,
               std::vector<int> &last_call)
{
	std::vector<std::pair<size_t, size_t>> todo;
	todo.emplace_back((size_t)0, members.size());
	std::string err;
	while (!todo.empty()) {
		const std::pair<size_t, size_t> r = todo.back();
		todo.pop_back();
		if (r.first >= r.second) continue;
		if (run_range(S, d, prm, reads, members, r.first, r.second, last_call, err)) continue;
		if (r.second - r.first == 1) {
			ReadRef &rd = reads[(size_t)members[r.first]];
			rd.err = true; rd.n_rec = rd.n_mini = 0; rd.in_arena = false;
			S.n_err.fetch_add(1);
			fprintf(stderr, "[mm2chain-b200] driver shim: read %u goes back to the host (err_flag = 1): %s\n", rd.task.read_id, err.c_str());
			continue;
		}
		const size_t mid = r.first + (r.second - r.first) / 2;
		todo.emplace_back(mid, r.second);
		todo.emplace_back(r.first, mid);
	}
}

// The whole batch in one device call with no host copy of a packet (seed_internal.h: seed_packets_begin / _end): the request
// packets are DMA'd as they are, the response packets are laid out in HBM and DMA'd into their buffers.  false = the device
// refused the batch as a whole (the caller then goes the careful way: halves, err_flag for a single refused read).
bool process_batch_direct(Shim &S, ShimDevice &d, const mm2b200_seed_params_t &prm, const std::vector<Packet> &pkts, const std::vector<PktHead> &heads,
                          const std::vector<std::pair<size_t, size_t>> &span, const std::vector<ReadRef> &reads, std::vector<Packet> &out)
{
	const size_t P = pkts.size(), R = reads.size();
	static const bool trace = [] { const char *e = getenv("MM2B200_TRACE"); return e && atoi(e); }();
	struct timespec t0, t1, t2, t3;
	if (trace) clock_gettime(CLOCK_MONOTONIC, &t0);
	std::vector<const void*> req(P);
	std::vector<int64_t> req_size(P), base(P);
	int64_t at = 0;
	for (size_t p = 0; p < P; ++p) { req[p] = pkts[p].buf; req_size[p] = pkts[p].size; base[p] = at; at += pkt_up64(pkts[p].size); }
	std::vector<int64_t> src16(R);
	std::vector<int32_t> n_mv(R), qlen(R);
	std::vector<uint32_t> bid(R);
	for (size_t k = 0; k < R; ++k) {
		const ReadRef &r = reads[k];
		src16[k] = (base[(size_t)r.pkt] + (r.mv - (const char*)pkts[(size_t)r.pkt].buf)) / 16;
		n_mv[k] = r.task.seednum; bid[k] = r.task.bid; qlen[k] = r.task.qlensum;
	}
	PacketCounts cnt;
	int rc = seed_packets_begin(d.idx, &prm, (int)P, req.data(), req_size.data(), (int64_t)R, src16.data(), n_mv.data(), bid.data(), qlen.data(), &cnt);
	if (rc != MM2B200_OK) {
		if ((rc == MM2B200_ERR_CUDA || rc == MM2B200_ERR_STATE) && shim_abort_on_error()) shim_fatal("device batch failed", mm2b200_last_error());
		return false;
	}
	if (trace) clock_gettime(CLOCK_MONOTONIC, &t1);
	// response layout (map.c:495-501,533-565): chaindp_sndhdr_t, then per read collect_result_t + records + mini_pos
	std::vector<int64_t> resp_size(P), rec_dst(R), mini_dst(R), hdr_dst;
	std::vector<char> hdr;
	hdr_dst.reserve(P + R); hdr.reserve(64 * (P + R));
	int64_t o = 0;
	for (size_t p = 0; p < P; ++p) {
		const int64_t start = o;
		PktHead oh;
		memset(&oh, 0, sizeof(oh));
		oh.magic = heads[p].magic; oh.tid = heads[p].tid; oh.num = heads[p].num; oh.type = heads[p].type; oh.lat = heads[p].lat;
		const size_t head_at = hdr.size();
		hdr_dst.push_back(o); hdr.insert(hdr.end(), (const char*)&oh, (const char*)&oh + sizeof(oh));
		o += sizeof(PktHead);
		for (size_t k = span[p].first; k < span[p].second; ++k) {
			const int64_t n_rec = cnt.rec_off[k + 1] - cnt.rec_off[k], n_mini = cnt.mini_off[k + 1] - cnt.mini_off[k];
			const int64_t rec_b = pkt_up64(n_rec * (int64_t)sizeof(struct new_seed)), mini_b = pkt_up64(n_mini * 8);
			PktResult pr;
			memset(&pr, 0, sizeof(pr));
			pr.read_id = reads[k].task.read_id;
			pr.n_a = (uint32_t)n_rec; pr.n_minipos = (uint32_t)n_mini; pr.rep_len = (uint32_t)cnt.rep_len[k];
			pr.sub_size = (uint32_t)(sizeof(PktResult) + rec_b + mini_b);
			hdr_dst.push_back(o); hdr.insert(hdr.end(), (const char*)&pr, (const char*)&pr + sizeof(pr));
			o += sizeof(PktResult);
			rec_dst[k] = o; o += rec_b;
			mini_dst[k] = o; o += mini_b;
		}
		resp_size[p] = o - start;
		if (resp_size[p] > 0x7fffffffLL) shim_fatal("response", "a response packet exceeds 2 GiB (fpga_get_retbuf reports its size as int)");
		const uint32_t sz = (uint32_t)resp_size[p];
		memcpy(hdr.data() + head_at + offsetof(PktHead, size), &sz, sizeof(sz));
	}
	std::vector<void*> resp(P);
	for (size_t p = 0; p < P; ++p) {
		resp[p] = g_pool.get((size_t)resp_size[p]);
		if (!resp[p]) shim_fatal("response", "out of host memory");
	}
	if (trace) clock_gettime(CLOCK_MONOTONIC, &t2);
	rc = seed_packets_end(d.idx, (int)P, resp.data(), resp_size.data(), (int64_t)R, rec_dst.data(), mini_dst.data(), (int64_t)hdr_dst.size(), hdr.data(), hdr_dst.data());
	if (rc != MM2B200_OK) {
		for (void *q : resp) g_pool.put(q);
		if ((rc == MM2B200_ERR_CUDA || rc == MM2B200_ERR_STATE) && shim_abort_on_error()) shim_fatal("device batch failed", mm2b200_last_error());
		return false;
	}
	if (trace) {
		clock_gettime(CLOCK_MONOTONIC, &t3);
		auto ms = [](const struct timespec &a, const struct timespec &b) { return 1e3 * (double)(b.tv_sec - a.tv_sec) + 1e-6 * (double)(b.tv_nsec - a.tv_nsec); };
		int64_t in_b = 0, out_b = 0;
		for (size_t p = 0; p < P; ++p) { in_b += pkts[p].size; out_b += resp_size[p]; }
		fprintf(stderr, "[mm2b200 trace] shim batch: %zu packets, %zu reads, %.1f MB in, %.1f MB out: upload + seeding + fill %.2f ms, layout + buffers %.2f ms, scatter + download %.2f ms\n",
		        P, R, 1e-6 * (double)in_b, 1e-6 * (double)out_b, ms(t0, t1), ms(t1, t2), ms(t2, t3));
	}
	S.n_batches.fetch_add(1);
	for (size_t p = 0; p < P; ++p) out.push_back(Packet{resp[p], (uint32_t)resp_size[p]});
	return true;
}

// parse, run, answer.  Appends one response packet per request packet to `out`.
void process_batch(Shim &S, ShimDevice &d, const mm2b200_seed_params_t &base, const std::vector<Packet> &pkts, std::vector<Packet> &out)
{
	std::vector<ReadRef> reads;
	std::vector<PktHead> heads(pkts.size());
	std::vector<std::pair<size_t, size_t>> span(pkts.size());      // reads[first, last) of packet p
	for (size_t p = 0; p < pkts.size(); ++p) {
		const char *b = (const char*)pkts[p].buf, *end = b + pkts[p].size;
		if (pkts[p].size < sizeof(PktHead)) shim_fatal("malformed request", "packet shorter than its header");
		memcpy(&heads[p], b, sizeof(PktHead));
		if (heads[p].type != MM2B200_PKT_TYPE) shim_fatal("malformed request", "chaindp_sndhdr_t.type is not 3 (map.c:301,491)");
		b += sizeof(PktHead);
		span[p].first = reads.size();
		for (int i = 0; i < (int)heads[p].num; ++i) {
			if (b + sizeof(PktTask) > end) shim_fatal("malformed request", "packet truncated at a task header");
			ReadRef r;
			r.pkt = (int)p; r.slot = i;
			memcpy(&r.task, b, sizeof(PktTask));
			if (r.task.seednum < 0) shim_fatal("malformed request", "negative minimizer count");
			r.mv = b + sizeof(PktTask);
			if (r.mv + (int64_t)r.task.seednum * 16 > end) shim_fatal("malformed request", "packet truncated inside a task");
			b = r.mv + pkt_up64((int64_t)r.task.seednum * 16);
			reads.push_back(std::move(r));
		}
		span[p].second = reads.size();
	}
	S.n_packets.fetch_add((int64_t)pkts.size());
	S.n_reads.fetch_add((int64_t)reads.size());
	// groups of equal (gap_ref, gap_qry): one in practice (map.c:358-366 derives both from the options for long reads)
	std::vector<std::pair<int32_t, int32_t>> keys;
	std::vector<std::vector<int>> groups;
	for (size_t k = 0; k < reads.size(); ++k) {
		const int32_t g_ref = reads[k].task.gap_ref, g_qry = reads[k].task.gap_qry;
		const std::pair<int32_t, int32_t> key(g_ref, g_qry);
		size_t g = 0;
		while (g < keys.size() && keys[g] != key) ++g;
		if (g == keys.size()) { keys.push_back(key); groups.emplace_back(); }
		groups[g].push_back((int)k);
	}
	if (S.direct && groups.size() == 1 && !reads.empty()) {
		mm2b200_seed_params_t q = base;
		q.chain.max_dist_x = keys[0].first; q.chain.max_dist_y = keys[0].second; q.chain.n_segs = keys[0].second;   // as below
		if (process_batch_direct(S, d, q, pkts, heads, span, reads, out)) return;
	}
	std::vector<int> last_call;
	for (size_t g = 0; g < groups.size(); ++g) {
		mm2b200_seed_params_t q = base;
		q.chain.max_dist_x = keys[g].first;       // max_chain_gap_ref, map.c:515
		q.chain.max_dist_y = keys[g].second;      // max_chain_gap_qry, map.c:516
		q.chain.n_segs = keys[g].second;          // the fork's device model reads n_segs from gap_qry (map.c:517)
		run_group(S, d, q, reads, groups[g], last_call);
	}
	// response packets (map.c:495-501,533-565); padding is zero
	for (size_t p = 0; p < pkts.size(); ++p) {
		int64_t need = sizeof(PktHead);
		for (size_t k = span[p].first; k < span[p].second; ++k) {
			const ReadRef &r = reads[k];
			need += (int64_t)sizeof(PktResult);
			if (!r.err) need += pkt_up64(r.n_rec * (int64_t)sizeof(struct new_seed)) + pkt_up64(r.n_mini * 8);
		}
		if (need > 0x7fffffffLL) shim_fatal("response", "a response packet exceeds 2 GiB (fpga_get_retbuf reports its size as int)");
		char *o = (char*)g_pool.get((size_t)need);
		if (!o) shim_fatal("response", "out of host memory");
		PktHead oh;
		memset(&oh, 0, sizeof(oh));
		oh.magic = heads[p].magic; oh.tid = heads[p].tid; oh.num = heads[p].num; oh.type = heads[p].type; oh.lat = heads[p].lat;
		oh.size = (uint32_t)need;
		memcpy(o, &oh, sizeof(oh));
		char *w = o + sizeof(oh);
		for (size_t k = span[p].first; k < span[p].second; ++k) {
			const ReadRef &r = reads[k];
			PktResult pr;
			memset(&pr, 0, sizeof(pr));
			pr.read_id = r.task.read_id;
			if (r.err) {
				pr.err_flag = 1; pr.sub_size = sizeof(PktResult);
				memcpy(w, &pr, sizeof(pr));
				w += sizeof(pr);
				continue;
			}
			const int64_t rec_b = pkt_up64(r.n_rec * (int64_t)sizeof(struct new_seed)), mini_b = pkt_up64(r.n_mini * 8);
			pr.n_a = (uint32_t)r.n_rec; pr.n_minipos = (uint32_t)r.n_mini; pr.rep_len = (uint32_t)r.rep_len;
			pr.sub_size = (uint32_t)(sizeof(PktResult) + rec_b + mini_b);
			memcpy(w, &pr, sizeof(pr));
			w += sizeof(pr);
			const size_t rec_n = (size_t)r.n_rec * sizeof(struct new_seed), mini_n = (size_t)r.n_mini * 8;
			if (rec_n) memcpy(w, r.rec, rec_n);
			memset(w + rec_n, 0, (size_t)rec_b - rec_n);               // padding to 64 bytes is zero (the buffer is recycled)
			w += rec_b;
			if (mini_n) memcpy(w, r.mini, mini_n);
			memset(w + mini_n, 0, (size_t)mini_b - mini_n);
			w += mini_b;
		}
		out.push_back(Packet{o, (uint32_t)need});
	}
}

void worker_main(Shim *S, ShimDevice *d)
{
	std::vector<Packet> take, out;
	for (;;) {
		bool commit = false;
		mm2b200_seed_params_t prm;
		{
			std::unique_lock<std::mutex> lk(S->mu);
			S->cv_work.wait(lk, [&] { return S->stop || !S->pending.empty(); });
			if (S->pending.empty()) break;       // stop requested and nothing left
			take.clear();
			int64_t mv = 0;
			while (!S->pending.empty() && (take.empty() || mv < S->batch_mv)) {
				take.push_back(S->pending.front());
				mv += S->pending.front().size / 16;
				S->pending.pop_front();
			}
			++S->busy;
			commit = d->index_dirty; d->index_dirty = false;
			prm = S->prm;
			if (!S->have_params) shim_fatal("request before fpga_set_params", "the DP parameters are not in the packets (main.c:243)");
			if (!S->have_index) shim_fatal("request before fpga_load_index", "no index streams were loaded (main.c:201-204)");
		}
		if (commit && mm2b200_index_commit(d->idx) != MM2B200_OK) shim_fatal("index commit failed", mm2b200_last_error());
		out.clear();
		process_batch(*S, *d, prm, take, out);
		int64_t bytes = 0;
		for (const Packet &p : take) { bytes += p.size; g_pool.put(p.buf); }
		{
			std::lock_guard<std::mutex> lk(S->mu);
			for (const Packet &p : out) S->done.push_back(p);
			S->pending_bytes -= bytes;
			--S->busy;
			S->cv_ret.notify_all();
			if (S->pending.empty() && S->busy == 0) S->cv_idle.notify_all();
		}
	}
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------------
// fpga.h
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int fpga_init(int flag)
{
	(void)flag;
	std::lock_guard<std::mutex> gl(g_shim_mu);
	if (g_shim) return 0;
	std::unique_ptr<Shim> S(new Shim());
	S->max_bytes = env_i64("MM2B200_SHIM_MAX_BYTES", S->max_bytes);
	S->batch_mv = env_i64("MM2B200_SHIM_BATCH_MV", S->batch_mv);
	{ const char *e = getenv("MM2B200_SHIM_ZEROCOPY"); S->direct = !(e && e[0] == '0'); }
	{ const std::vector<int> devs = shim_devices(); g_pool.pin = S->direct; g_pool.device = devs.empty() ? 0 : devs[0]; }
	// two workers per GPU by default (each with its own context and copy of the index streams): while one batch is on the
	// device the other worker parses packets / lays out responses (and, on the careful path, copies them)
	const int per_dev = (int)env_i64("MM2B200_SHIM_WORKERS", 2);
	for (int ord : shim_devices()) for (int rep = 0; rep < per_dev; ++rep) {
		std::unique_ptr<ShimDevice> d(new ShimDevice());
		d->ordinal = ord;
		mm2b200_config_t cfg;
		memset(&cfg, 0, sizeof(cfg));
		cfg.device = ord;
		cfg.n_slots = 1;
		if (mm2b200_create(&cfg, &d->ctx) != MM2B200_OK || mm2b200_index_create(d->ctx, &d->idx) != MM2B200_OK) {
			fprintf(stderr, "[mm2chain-b200] fpga_init: device %d: %s\n", ord, mm2b200_last_error());
			if (shim_abort_on_error()) abort();
			return -1;
		}
		S->dev.push_back(std::move(d));
	}
	for (auto &d : S->dev) d->th = std::thread(worker_main, S.get(), d.get());
	g_shim = S.release();
	return 0;
}

extern "C" void fpga_finalize()
{
	std::lock_guard<std::mutex> gl(g_shim_mu);
	Shim *S = g_shim;
	if (!S) return;
	{ std::lock_guard<std::mutex> lk(S->mu); S->stop = true; }
	S->cv_work.notify_all();
	for (auto &d : S->dev) if (d->th.joinable()) d->th.join();
	for (auto &d : S->dev) {
		if (d->mv) mm2b200_pinned_free(d->mv);
		if (d->idx) mm2b200_index_destroy(d->idx);
		if (d->ctx) mm2b200_destroy(d->ctx);
	}
	for (const Packet &p : S->done) g_pool.put(p.buf);
	for (const Packet &p : S->pending) g_pool.put(p.buf);
	g_pool.drain();
	g_shim = nullptr;
	delete S;
}

extern "C" void fpga_set_params(int bw, int is_cdna, int max_skip, int min_sc, int flag, int max_occ)
{
	Shim *S = g_shim;
	if (!S) shim_fatal("fpga_set_params", "fpga_init has not run");
	std::lock_guard<std::mutex> lk(S->mu);
	memset(&S->prm, 0, sizeof(S->prm));
	S->prm.chain.bw = bw; S->prm.chain.is_cdna = is_cdna; S->prm.chain.max_skip = max_skip; S->prm.chain.min_sc = min_sc;
	S->prm.chain.min_cnt = 1;                  // stays on the host (mm_chain_dp_bottom); unused by the device half
	S->prm.chain.n_segs = 1; S->prm.chain.max_iter = 0;
	S->prm.flag = flag; S->prm.max_occ = max_occ;
	S->have_params = true;
	S->index_open = false;
}

extern "C" void fpga_load_index(void *addr, int size, int type)
{
	Shim *S = g_shim;
	if (!S) shim_fatal("fpga_load_index", "fpga_init has not run");
	std::unique_lock<std::mutex> lk(S->mu);
	// a new index part replaces the old one: everything submitted against the old one must be through
	S->cv_idle.wait(lk, [&] { return S->pending.empty() && S->busy == 0; });
	if (!S->index_open) {              // the first chunk of a new index: nothing of the old one may linger, served or not
		for (auto &d : S->dev) if (mm2b200_index_reset(d->idx) != MM2B200_OK) shim_fatal("fpga_load_index", mm2b200_last_error());
		S->index_open = true;
	}
	for (auto &d : S->dev) {
		if (mm2b200_index_load(d->idx, addr, size, type) != MM2B200_OK) shim_fatal("fpga_load_index", mm2b200_last_error());
		d->index_dirty = true;
	}
	S->have_index = true;
}

extern "C" void *fpga_get_writebuf_thread(unsigned long size, BUF_TYPE type, int tid)
{
	(void)type; (void)tid;
	Shim *S = g_shim;
	if (!S) return nullptr;
	{
		std::lock_guard<std::mutex> lk(S->mu);
		if (S->pending_bytes > 0 && S->pending_bytes + (int64_t)size > S->max_bytes) return nullptr;   // back-pressure: the caller sleeps and retries
		S->pending_bytes += (int64_t)size;
	}
	void *p = g_pool.get(size ? size : 1);
	if (!p) { std::lock_guard<std::mutex> lk(S->mu); S->pending_bytes -= (int64_t)size; }
	return p;
}

extern "C" void *fpga_get_writebuf(unsigned long size, BUF_TYPE type) { return fpga_get_writebuf_thread(size, type, -1); }

extern "C" int fpga_writebuf_submit(void *addr, unsigned int size, unsigned int type)
{
	(void)type;
	Shim *S = g_shim;
	if (!S || !addr) return -1;
	{
		std::lock_guard<std::mutex> lk(S->mu);
		S->index_open = false;
		S->pending.push_back(Packet{addr, size});
	}
	S->cv_work.notify_one();
	return 0;
}

extern "C" void *fpga_get_retbuf(int *len, RET_TYPE type)
{
	(void)type;
	Shim *S = g_shim;
	if (len) *len = 0;
	if (!S) return nullptr;
	std::unique_lock<std::mutex> lk(S->mu);
	S->cv_ret.wait(lk, [&] { return !S->done.empty() || (S->exit_block && S->pending.empty() && S->busy == 0); });
	if (S->done.empty()) return nullptr;
	const Packet p = S->done.front();
	S->done.pop_front();
	if (len) *len = (int)p.size;
	return p.buf;
}

extern "C" int fpga_release_retbuf(void *addr) { g_pool.put(addr); return 0; }

extern "C" void fpga_exit_block()
{
	Shim *S = g_shim;
	if (!S) return;
	{ std::lock_guard<std::mutex> lk(S->mu); S->exit_block = true; }
	S->cv_ret.notify_all();
}

extern "C" void fpga_set_block()
{
	Shim *S = g_shim;
	if (!S) return;
	std::lock_guard<std::mutex> lk(S->mu);
	S->exit_block = false;
}

extern "C" int mm2b200_shim_stats(int64_t *n_packets, int64_t *n_reads, int64_t *n_batches, int64_t *n_err_reads, int64_t *n_launches)
{
	Shim *S = g_shim;
	if (!S) return -1;
	if (n_packets) *n_packets = S->n_packets.load();
	if (n_reads) *n_reads = S->n_reads.load();
	if (n_batches) *n_batches = S->n_batches.load();
	if (n_err_reads) *n_err_reads = S->n_err.load();
	if (n_launches) {
		int64_t n = 0;
		for (auto &d : S->dev) n += mm2b200_launch_count(d->ctx) + mm2b200_seed_launch_count(d->idx);
		*n_launches = n;
	}
	return 0;
}
