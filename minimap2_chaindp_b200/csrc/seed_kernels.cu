// seed_kernels.cu -- seed-hit collection on the device: mm_idx_get over the fork's index
// streams, collect_matches / skip_seed / collect_seed_hits (map.c:112-236) and the per-read sort (map.c:233), producing the
// sorted anchor arrays the chaining kernels read -- in HBM, never on the host.
//
// Three stages:
//   lookup   one warp per read, 32 minimizers per step: hash probe (random 16 B + 64 B + 8 B reads), occurrence counts,
//            the repetitive-interval bookkeeping (rep_len), mini_pos[], and the offset of every minimizer's first anchor
//            within its read; then exclusive scans over the reads (anchor offsets, mini_pos offsets).
//   expand   one thread per minimizer: reads its positions, applies skip_seed, writes anchors in the reference's
//            collection order plus a 64-bit sort key per anchor.
//   sort     one CTA per read: keys sorted in shared memory (stable by construction: the anchor's number is the low part
//            of the key).  radix_sort_128x is NOT stable beyond 64 elements (ksort.h:116-151), so reads that have equal
//            x values and more than 64 anchors are re-sorted by a step-for-step replay of the reference's sort.
#include <cuda_runtime.h>
#include <stdint.h>
#include "seed_internal.h"
#include "sort_exact.cuh"

namespace mm2b200 {

#define F_NO_DIAG   0x001      // minimap.h:8
#define F_NO_DUAL   0x002      // minimap.h:9
#define F_FOR_ONLY  0x100000   // minimap.h:28
#define F_REV_ONLY  0x200000   // minimap.h:29
#define P_STRAND_BIT 21        // MM_P_STRAND, mmpriv.h:20
#define SEED_TANDEM (1ULL << 42)   // mmpriv.h:18
#define SEED_SELF   (1ULL << 43)   // mmpriv.h:19
#define SEG_SHIFT   48             // mmpriv.h:22

#define CNT_SINGLE (1 << 30)
#define CNT_TANDEM (1u << 31)
#define CNT_MASK   ((1 << 30) - 1)

struct Hit { int t; int single; uint64_t v; };

// mm_idx_get (index.c:221-238) with khash's probe sequence (khash.h:218-231; hash = key>>1, index.c:23)
static __device__ __forceinline__ Hit idx_get(const DevIndex &ix, uint64_t minier, int64_t *err)
{
	Hit h; h.t = 0; h.single = 0; h.v = 0;
	if (ix.bloom_mask) {                              // absent for sure when either bit is clear
		const uint64_t b0 = bloom_bit(minier, 0, ix.bloom_mask), b1 = bloom_bit(minier, 1, ix.bloom_mask);
		if (!((__ldg(ix.bloom + (b0 >> 5)) >> (b0 & 31)) & 1u) || !((__ldg(ix.bloom + (b1 >> 5)) >> (b1 & 31)) & 1u)) return h;
	}
	const uint4 rec = __ldg(ix.B + (minier & ((1ULL << ix.b) - 1)));
	const uint64_t b0 = (uint64_t)rec.y << 32 | rec.x, b1 = (uint64_t)rec.w << 32 | rec.z;
	const uint64_t h_off = b1 >> 28, p_off = (b1 & 0xfffffffULL) << 8 | b0 >> 56;
	const uint32_t n_buckets = (uint32_t)(b0 >> 24);
	if (n_buckets == 0) return h;
	const uint64_t want = minier >> ix.b;            // key >> 1
	const uint32_t m = n_buckets - 1;
	uint32_t i = (uint32_t)want & m, step = 0;
	const uint32_t last = i;
	for (;;) {
		const uint64_t slot = h_off + i;
		if (slot >= (uint64_t)ix.n_slots) { atomicOr((unsigned long long*)err, (unsigned long long)SE_INDEX_RANGE); return h; }
		const uint8_t *grp = ix.H + 64 * (slot >> 3);
		const uint32_t fl = (__ldg((const uint32_t*)grp) >> ((i & 0xfU) << 1)) & 3;
		if (fl & 2) return h;                        // empty slot: absent
		const uint16_t *kp = (const uint16_t*)(grp + 4 + 6 * (slot & 7));
		const uint64_t kk = (uint64_t)__ldg(kp) | (uint64_t)__ldg(kp + 1) << 16 | (uint64_t)__ldg(kp + 2) << 32;
		if (!(fl & 1) && (kk >> 1) == want) {
			const uint64_t v = __ldg(ix.V + slot);
			if (kk & 1) { h.t = 1; h.single = 1; h.v = v; return h; }      // index.c:231-233
			h.t = (int)(uint32_t)v; h.v = p_off + (v >> 32);                // index.c:235-236
			if (h.t < 0 || h.v + (uint64_t)h.t > (uint64_t)ix.n_pos) { atomicOr((unsigned long long*)err, (unsigned long long)SE_INDEX_RANGE); h.t = 0; }
			return h;
		}
		i = (i + (++step)) & m;
		if (i == last) return h;
	}
}

// one CTA per bucket: every live slot of its hash table sets the key's two bits
__global__ void __launch_bounds__(256) k_seed_bloom(DevIndex ix, uint32_t *bloom, int32_t *bad)
{
	const int64_t n_bkt = (int64_t)1 << ix.b;
	for (int64_t bk = blockIdx.x; bk < n_bkt; bk += gridDim.x) {
		const uint4 rec = ix.B[bk];
		const uint64_t b0 = (uint64_t)rec.y << 32 | rec.x, b1 = (uint64_t)rec.w << 32 | rec.z;
		const uint64_t h_off = b1 >> 28;
		const uint32_t n_buckets = (uint32_t)(b0 >> 24);
		for (uint32_t i = threadIdx.x; i < n_buckets; i += blockDim.x) {
			const uint64_t slot = h_off + i;
			if (slot >= (uint64_t)ix.n_slots) { *bad = 1; break; }    // inconsistent streams: no filter, the lookup reports them
			const uint8_t *grp = ix.H + 64 * (slot >> 3);
			const uint32_t fl = (*(const uint32_t*)grp >> ((i & 0xfU) << 1)) & 3;
			if (fl) continue;                                        // empty or deleted
			const uint16_t *kp = (const uint16_t*)(grp + 4 + 6 * (slot & 7));
			const uint64_t kk = (uint64_t)kp[0] | (uint64_t)kp[1] << 16 | (uint64_t)kp[2] << 32;
			const uint64_t minier = (kk >> 1) << ix.b | (uint64_t)bk;
			for (int wch = 0; wch < 2; ++wch) {
				const uint64_t bit = bloom_bit(minier, wch, ix.bloom_mask);
				atomicOr(bloom + (bit >> 5), 1u << (bit & 31));
			}
		}
	}
}

int launch_seed_bloom(const DevIndex &ix, uint32_t *bloom, int32_t *bad, cudaStream_t s)
{
	const int64_t n_bkt = (int64_t)1 << ix.b;
	k_seed_bloom<<<(int)(n_bkt < 65535 ? n_bkt : 65535), 256, 0, s>>>(ix, bloom, bad);
	return 1;
}

// skip_seed (map.c:146-185).  Only bit 0 of flag opens the name-order block ("1 & flag & (...)", map.c:150).
static __device__ __forceinline__ bool skip_seed(int flag, uint64_t r, uint32_t q_pos, uint32_t bid, bool *is_self)
{
	*is_self = false;
	if (1 & flag & (F_NO_DIAG | F_NO_DUAL)) {
		const uint32_t rank = (uint32_t)r & 0x1FFFFF, flg = bid >> 31, val = bid & 0x7fffffff;
		const int cmp = val > rank ? 1 : val < rank ? -1 : flg ? 0 : -1;
		if ((flag & F_NO_DIAG) && cmp == 0) {
			if (((r >> 22) & 0x1fffff) == (q_pos >> 1)) return true;
			if (((r >> P_STRAND_BIT) & 1) == (q_pos & 1)) *is_self = true;
		}
		if ((flag & F_NO_DUAL) && cmp > 0) return true;
	}
	if (flag & (F_FOR_ONLY | F_REV_ONLY)) {
		if (((r >> P_STRAND_BIT) & 1) == (q_pos & 1)) { if (flag & F_REV_ONLY) return true; }
		else if (flag & F_FOR_ONLY) return true;
	}
	return false;
}

static __device__ __forceinline__ bool filters_on(int flag) { return (1 & flag & (F_NO_DIAG | F_NO_DUAL)) || (flag & (F_FOR_ONLY | F_REV_ONLY)); }

// ---------------------------------------------------------------------------------------------------------
// k_seed_lookup: collect_matches (map.c:112-144) for one read per warp
// ---------------------------------------------------------------------------------------------------------
#define LOOKUP_T 256

__global__ void __launch_bounds__(LOOKUP_T, 6) k_seed_lookup(DevIndex ix, SeedWork w)
{
	const int lane = threadIdx.x & 31;
	const unsigned lt = (1u << lane) - 1;
	const bool filt = filters_on(w.flag);
	// reads are handed out one by one through a counter (the grid is what is resident; reads differ in length and in how many
	// of their minimizers are in the index)
	for (;;) {
		int r = 0;
		if (lane == 0) r = (int)atomicAdd((unsigned long long*)(w.totals + ST_NEXT_READ), 1ULL);
		r = __shfl_sync(0xffffffffu, r, 0);
		if (r >= w.n_reads) break;
		const int64_t m0 = w.mv_off[r], m1 = w.mv_off[r + 1];
		const uint32_t bid = w.bid[r];
		int run_a = 0, run_m = 0, rep_sum = 0, carry_en = 0;     // carry_en: rep_en of map.c:114 (rep_st = rep_en = 0 at the start)
		for (int64_t base = m0; base < m1; base += 32) {
			const int64_t i = base + lane;
			const bool valid = i < m1;
			ulonglong2 mz = make_ulonglong2(0, 0);
			if (valid) mz = w.mv[i];
			const uint32_t q_pos = (uint32_t)mz.y, q_span = (uint32_t)(mz.x & 0xff);
			Hit h; h.t = 0; h.single = 0; h.v = 0;
			if (valid) h = idx_get(ix, mz.x >> 8, w.totals + ST_ERROR);
			const bool rep = valid && h.t >= w.max_occ;           // map.c:126
			const bool keep = valid && !rep;
			// repetitive intervals (map.c:127-131): an interval opens where st > rep_en of the previous repetitive minimizer
			const int en = (int)(q_pos >> 1) + 1, st = en - (int)q_span;
			const unsigned rep_b = __ballot_sync(0xffffffffu, rep);
			const unsigned before = rep_b & lt;
			const int pl = before ? 31 - __clz(before) : 0;
			int pe = __shfl_sync(0xffffffffu, en, pl);
			if (!before) pe = carry_en;
			int contrib = (rep && st > pe) ? pe - st : 0;          // closes the previous interval at pe, opens one at st
			contrib = __reduce_add_sync(0xffffffffu, contrib);
			rep_sum += contrib;
			if (rep_b) carry_en = __shfl_sync(0xffffffffu, en, 31 - __clz(rep_b));
			// anchors this minimizer contributes
			int kept = 0;
			uint32_t tandem = 0;
			if (keep) {
				if (i > m0 && (w.mv[i - 1].x >> 8) == (mz.x >> 8)) tandem = CNT_TANDEM;        // map.c:135-137
				if (i < m1 - 1 && (w.mv[i + 1].x >> 8) == (mz.x >> 8)) tandem = CNT_TANDEM;
				if (!filt) kept = h.t;
				else {
					for (int k = 0; k < h.t; ++k) {
						const uint64_t pos = h.single ? h.v : __ldg(ix.P + h.v + k);
						bool self;
						if (!skip_seed(w.flag, pos, q_pos, bid, &self)) ++kept;
					}
				}
			}
			int incl = kept;
			#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
			if (valid) {
				w.src[i] = h.v;
				w.cnt[i] = keep ? (int32_t)((uint32_t)h.t | (h.single ? CNT_SINGLE : 0) | tandem) : 0;
				w.aofs[i] = run_a + incl - kept;
				w.mread[i] = r;
			}
			run_a += __shfl_sync(0xffffffffu, incl, 31);
			// mini_pos (map.c:139): every minimizer that is not repetitive, found or not
			const unsigned keep_b = __ballot_sync(0xffffffffu, keep);
			if (keep) w.mini_sparse[m0 + run_m + __popc(keep_b & lt)] = (uint64_t)q_span << 32 | q_pos >> 1;
			run_m += __popc(keep_b);
		}
		if (lane == 0) {
			w.n_a[r] = run_a;
			w.n_mini[r] = run_m;
			w.rep_len[r] = rep_sum + carry_en;                     // map.c:142
		}
	}
}

// exclusive scans of n_a[] and n_mini[] over the reads by one CTA; totals
#define SCAN_T 1024
__global__ void __launch_bounds__(SCAN_T) k_seed_scan(SeedWork w)
{
	__shared__ long long s_a[SCAN_T], s_m[SCAN_T];
	__shared__ int s_max[SCAN_T];
	const int tid = threadIdx.x;
	const int per = (w.n_reads + SCAN_T - 1) / SCAN_T;
	const int r0 = min(tid * per, w.n_reads), r1 = min(r0 + per, w.n_reads);
	long long sa = 0, sm = 0; int mx = 0;
	for (int r = r0; r < r1; ++r) { const int n = w.n_a[r]; sa += n; sm += w.n_mini[r]; mx = max(mx, n); }
	s_a[tid] = sa; s_m[tid] = sm; s_max[tid] = mx;
	__syncthreads();
	for (int d = 1; d < SCAN_T; d <<= 1) {
		long long va = 0, vm = 0; int vx = 0;
		if (tid >= d) { va = s_a[tid - d]; vm = s_m[tid - d]; vx = s_max[tid - d]; }
		__syncthreads();
		s_a[tid] += va; s_m[tid] += vm; s_max[tid] = max(s_max[tid], vx);
		__syncthreads();
	}
	long long ea = s_a[tid] - sa, em = s_m[tid] - sm;
	for (int r = r0; r < r1; ++r) { w.a_off[r] = ea; w.mini_off[r] = em; ea += w.n_a[r]; em += w.n_mini[r]; }
	if (tid == SCAN_T - 1) {
		w.a_off[w.n_reads] = s_a[tid]; w.mini_off[w.n_reads] = s_m[tid];
		w.totals[ST_ANCHORS] = s_a[tid]; w.totals[ST_MINI] = s_m[tid]; w.totals[ST_MAX_READ] = s_max[tid];
		if (s_max[tid] > SEED_MAX_READ_ANCHORS) atomicOr((unsigned long long*)(w.totals + ST_ERROR), (unsigned long long)SE_READ_TOO_LONG);
	}
}

// mini_pos of all reads back to back
__global__ void k_seed_pack_mini(SeedWork w)
{
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= w.n_mv) return;
	const int r = w.mread[i];
	const int64_t k = i - w.mv_off[r];
	if (k < w.n_mini[r]) w.mini_out[w.mini_off[r] + k] = w.mini_sparse[i];
}

int launch_seed_lookup(const DevIndex &ix, const SeedWork &w, int sm_count, cudaStream_t s)
{
	cudaMemsetAsync(w.totals, 0, ST_TOTALS * sizeof(int64_t), s);
	int n = 0;
	if (w.n_reads > 0 && w.n_mv > 0) {
		const int warps_per_cta = LOOKUP_T / 32;
		int grid = (w.n_reads + warps_per_cta - 1) / warps_per_cta;
		// resident CTAs per SM (the same on every device this process uses; initialised once, also when several shim workers get here together)
		static const int occ = [] { int o = 0; if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, k_seed_lookup, LOOKUP_T, 0) != cudaSuccess || o < 1) { cudaGetLastError(); o = 4; } return o; }();
		const int cap = sm_count * occ;               // one wave: a grid beyond it would run its rest on a part of the machine
		if (grid > cap) grid = cap;
		k_seed_lookup<<<grid, LOOKUP_T, 0, s>>>(ix, w); ++n;
	} else if (w.n_reads > 0) {
		cudaMemsetAsync(w.n_a, 0, 4 * (size_t)w.n_reads, s); cudaMemsetAsync(w.n_mini, 0, 4 * (size_t)w.n_reads, s);
		cudaMemsetAsync(w.rep_len, 0, 4 * (size_t)w.n_reads, s);
	}
	k_seed_scan<<<1, SCAN_T, 0, s>>>(w); ++n;
	if (w.n_mv > 0) { k_seed_pack_mini<<<(unsigned)((w.n_mv + 255) / 256), 256, 0, s>>>(w); ++n; }
	return n;
}

// ---------------------------------------------------------------------------------------------------------
// k_seed_expand: the anchor loop of collect_seed_hits (map.c:197-230), one thread per minimizer
// ---------------------------------------------------------------------------------------------------------
// order-preserving 43-bit image of an anchor's x (strand 1 | refid 21 | refpos 21; bits 21..31 and 53..62 of x are zero)
static __device__ __forceinline__ uint64_t squeeze_x(uint64_t x) { return (x >> 63) << 42 | ((x >> 32) & 0x1fffff) << 21 | (x & 0x1fffff); }

__global__ void __launch_bounds__(256) k_seed_expand(DevIndex ix, SeedWork w)
{
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= w.n_mv) return;
	const uint32_t c = (uint32_t)w.cnt[i];
	const int t = (int)(c & CNT_MASK);
	if (t == 0) return;
	const int r = w.mread[i];
	const ulonglong2 mz = w.mv[i];
	const uint32_t q_pos = (uint32_t)mz.y, q_span = (uint32_t)(mz.x & 0xff);
	const uint32_t bid = w.bid[r], qlen = (uint32_t)w.qlen[r];
	const uint64_t src = w.src[i];
	const uint64_t seg = ((mz.y >> 32) & 0x7fffffff) << SEG_SHIFT;       // 31-bit seg_id (map.c:107,134), then map.c:225
	const uint64_t y_fwd = (uint64_t)q_span << 32 | q_pos >> 1;          // map.c:219
	const uint64_t y_rev = (uint64_t)q_span << 32 | (uint32_t)(qlen - ((q_pos >> 1) + 1 - q_span) - 1);   // map.c:222, 32-bit unsigned like the reference
	const uint64_t y_add = seg | ((c & CNT_TANDEM) ? SEED_TANDEM : 0);
	int idx = w.aofs[i];
	const int64_t base = w.a_off[r];
	for (int k = 0; k < t; ++k) {
		const uint64_t pos = (c & CNT_SINGLE) ? src : __ldg(ix.P + src + k);
		bool self;
		if (skip_seed(w.flag, pos, q_pos, bid, &self)) continue;
		const uint64_t rpos = (pos >> 22) & 0x1fffff;
		const bool fwd = ((pos >> P_STRAND_BIT) & 1) == (q_pos & 1);
		ulonglong2 a;
		a.x = (fwd ? 0 : 1ULL << 63) | ((pos & 0xfffff80000000000ULL) >> 11) | rpos;     // map.c:218,221
		a.y = (fwd ? y_fwd : y_rev) | y_add | (self ? SEED_SELF : 0);
		w.a_un[base + idx] = a;
		w.key[base + idx] = squeeze_x(a.x) << SEED_IDX_BITS | (uint64_t)(idx & (SEED_MAX_READ_ANCHORS - 1));
		++idx;
	}
}

int launch_seed_expand(const DevIndex &ix, const SeedWork &w, cudaStream_t s)
{
	if (w.n_mv <= 0) return 0;
	k_seed_expand<<<(unsigned)((w.n_mv + 255) / 256), 256, 0, s>>>(ix, w);
	return 1;
}

// ---------------------------------------------------------------------------------------------------------
// k_seed_sort: radix_sort_128x(a, a + n_a) (map.c:233) for one read per CTA
// ---------------------------------------------------------------------------------------------------------
#define SORT_T 256
#define SORT_CAP 4096       // keys of one read held in shared memory (32 KB); longer reads sort in their global key array

template <int T>
static __device__ void cta_sort_keys(uint64_t *z, int n)
{	// "flip" bitonic network: every compare-exchange is ascending, so no padding to a power of two is needed
	const int tid = threadIdx.x;
	for (int k = 2; (k >> 1) < n; k <<= 1) {
		for (int i = tid; i < n; i += T) {
			const int l = i ^ (k - 1);
			if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
		}
		__syncthreads();
		for (int j = k >> 2; j > 0; j >>= 1) {
			for (int i = tid; i < n; i += T) {
				const int l = i ^ j;
				if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
			}
			__syncthreads();
		}
	}
}

// The same network for keys in shared memory, with the compare-exchanges at distances below 32 done in registers: a warp
// holds 32 consecutive keys, partners are exchanged with shuffles (no shared-memory round trip, no block barrier) -- 40 of
// the 55 passes a read of 513..1024 anchors needs.  Passes at distance >= 32 go through shared memory, one thread per PAIR.
// z[] holds n keys and room up to the next multiple of 32 (filled with the largest key, which stays behind the real ones).
static __device__ __forceinline__ uint64_t cx_lane(uint64_t v, int lane, int m)
{
	const uint64_t o = __shfl_xor_sync(0xffffffffu, v, m);
	const bool lower = lane < (lane ^ m);
	return lower == (v < o) ? v : o;                   // the lower lane keeps the smaller key
}

template <int T>
static __device__ void cta_sort_keys_sh(uint64_t *z, int n)
{
	const int tid = threadIdx.x, lane = tid & 31;
	const int np = (n + 31) & ~31;
	for (int i = n + tid; i < np; i += T) z[i] = ~0ULL;
	__syncthreads();
	for (int e = tid; e < np; e += T) {                 // blocks k = 2 .. 32: entirely inside a warp
		uint64_t v = z[e];
		#pragma unroll
		for (int k = 2; k <= 32; k <<= 1) {
			v = cx_lane(v, lane, k - 1);
			#pragma unroll
			for (int j = k >> 2; j > 0; j >>= 1) v = cx_lane(v, lane, j);
		}
		z[e] = v;
	}
	__syncthreads();
	for (int k = 64; (k >> 1) < np; k <<= 1) {
		const int half = k >> 1, hs = __ffs(half) - 1, n_pairs = ((np + k - 1) >> (hs + 1)) << hs;
		for (int p = tid; p < n_pairs; p += T) {        // the flip: i <-> i ^ (k - 1) inside blocks of k
			const int i = ((p >> hs) << (hs + 1)) + (p & (half - 1)), l = i ^ (k - 1);
			if (l < np) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
		}
		__syncthreads();
		for (int j = k >> 2; j >= 32; j >>= 1) {
			for (int p = tid; ; p += T) {               // pair p: i has bit j clear; i grows with p
				const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1)), l = i + j;
				if (l >= np) break;
				const uint64_t x = z[i], y = z[l];
				if (x > y) { z[i] = y; z[l] = x; }
			}
			__syncthreads();
		}
		for (int e = tid; e < np; e += T) {             // distances 16 .. 1
			uint64_t v = z[e];
			#pragma unroll
			for (int j = 16; j > 0; j >>= 1) v = cx_lane(v, lane, j);
			z[e] = v;
		}
		__syncthreads();
	}
}

__global__ void __launch_bounds__(SORT_T) k_seed_sort(SeedWork w)
{
	__shared__ uint64_t s_key[SORT_CAP];
	__shared__ int s_head[256], s_tail[256];
	__shared__ int s_tie;
	const int tid = threadIdx.x;
	for (int r = blockIdx.x; r < w.n_reads; r += gridDim.x) {
		const int64_t base = w.a_off[r];
		const int n = (int)(w.a_off[r + 1] - base);
		if (n == 0) continue;
		if (n > SEED_MAX_READ_ANCHORS) continue;          // flagged by k_seed_scan; the host fails the batch
		uint64_t *z = n <= SORT_CAP ? s_key : w.key + base;
		if (tid == 0) s_tie = 0;
		if (n <= SORT_CAP) for (int i = tid; i < n; i += SORT_T) s_key[i] = w.key[base + i];
		__syncthreads();
		if (n <= SORT_CAP) cta_sort_keys_sh<SORT_T>(z, n); else cta_sort_keys<SORT_T>(z, n);
		int tie = 0;
		for (int i = tid; i < n; i += SORT_T) {
			const uint64_t k = z[i];
			if (i + 1 < n && (z[i + 1] >> SEED_IDX_BITS) == (k >> SEED_IDX_BITS)) tie = 1;
			w.a[base + i] = w.a_un[base + (int)(k & (SEED_MAX_READ_ANCHORS - 1))];
		}
		if (n > RS_MIN && tie) s_tie = 1;
		__syncthreads();
		if (s_tie) {
			// equal keys in a read of more than 64 anchors: the reference's sort is not stable there -- replay it step for step
			// (in shared memory when both arrays fit where the sort keys were: one thread walks them with dependent accesses)
			const bool in_sh = n <= SORT_CAP / 2;
			uint64_t *kx = in_sh ? s_key : w.key + base, *ky = in_sh ? s_key + SORT_CAP / 2 : w.ky + base;
			for (int i = tid; i < n; i += SORT_T) { const ulonglong2 a = w.a_un[base + i]; kx[i] = a.x; ky[i] = a.y; }
			__syncthreads();
			if (tid == 0) {
				radix_sort_kv_exact(kx, ky, n, w.stk + base, s_head, s_tail);
				atomicAdd((unsigned long long*)(w.totals + ST_EXACT), 1ULL);
			}
			__syncthreads();
			for (int i = tid; i < n; i += SORT_T) w.a[base + i] = make_ulonglong2(kx[i], ky[i]);
		}
		__syncthreads();
	}
}

int launch_seed_sort(const SeedWork &w, cudaStream_t s)
{
	if (w.n_reads <= 0) return 0;
	const int grid = w.n_reads < (1 << 20) ? w.n_reads : (1 << 20);
	k_seed_sort<<<grid, SORT_T, 0, s>>>(w);
	return 1;
}

} // namespace mm2b200
