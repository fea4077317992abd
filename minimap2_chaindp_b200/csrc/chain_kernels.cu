// chain_kernels.cu -- sm_100a kernels for minimap2's anchor-chaining DP (reference: chain.c:23-216).
//
// Stage map (reference lines -> kernel):
//   chain.c:46-47  avg_qspan, run splitting          k_prep            warp per read
//   chain.c:52-90  DP fill f/p/v with the max_skip early exit       k_fill<GENERIC>   warp per run, persistent
//   chain.c:93-122 compaction to new_seed[] + new indices           k_compact         warp per read
//   chain.c:127-180 chain ends, peaks, ordering, backtrack, filters k_tail            CTA per read
//   chain.c:186-214 gather + first-anchor-x order                   k_tail* (order), k_scan_*, k_gather
//
// Everything is integer work on 16-byte anchors; nothing here is GEMM-shaped, so no tensor cores.  The one
// floating-point expression, (int)(dd * .01 * avg_qspan) (chain.c:73,78), is evaluated with explicit
// round-to-nearest double multiplies in the reference's association order so the result is bit-identical.
#include <cstdlib>
#include "chain_internal.h"
#include "sort_exact.cuh"

namespace mm2b200 {

#define FULL 0xffffffffu

static __device__ __forceinline__ unsigned lanemask_lt(int lane) { return (1u << lane) - 1u; }           // lane in [0,31]
static __device__ __forceinline__ unsigned mask_below(int n) { return n >= 32 ? FULL : ((1u << n) - 1u); } // n in [0,32]

// Run list: short runs are appended from the front, long ones (>= LONG_RUN anchors) from the back of units[], so the
// fill kernels can start the long runs first (longest-processing-time-first keeps the tail of the kernel short).
// The compiler aggregates the same-address atomics of a warp into one.
#define LONG_RUN 128
static __device__ __forceinline__ void emit_unit(const Workspace &w, int start, int len, int read)
{
	Unit u; u.start = start; u.len = len; u.read = read; u.pad = 0;
	if (len >= LONG_RUN) w.units[w.units_cap - 1 - atomicAdd(&w.counters[CNT_LONG_UNITS], 1)] = u;
	else w.units[atomicAdd(&w.counters[CNT_UNITS], 1)] = u;
}
// queue position q -> run: long runs first
static __device__ __forceinline__ Unit fetch_unit(const Workspace &w, int q, int n_long)
{
	return q < n_long ? w.units[w.units_cap - 1 - q] : w.units[q - n_long];
}

// ---------------------------------------------------------------------------------------------------------
// k_prep: per read -- avg_qspan (chain.c:46-47), singleton defaults for f/p/v, mark reset, and the list of
// independent runs (consecutive x gap <= max_dist_x; chain.c:58 makes st jump to the run start).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_prep(Workspace w, int32_t max_dist_x)
{
	const int lane = threadIdx.x & 31;
	const int gw = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
	const int nw = (int)((gridDim.x * (unsigned)blockDim.x) >> 5);
	const unsigned long long mdx = (unsigned long long)(long long)max_dist_x;
	for (int r = gw; r < w.n_reads; r += nw) {
		const int b = (int)w.offs[r], e = (int)w.offs[r + 1];
		unsigned long long sum = 0, prev_last_x = 0;
		int carry_start = b, multiseg = 0;
		// four 512-byte tiles are requested before the first is consumed (one tile in flight per warp left the kernel at
		// 44 % of the copy bandwidth); the tiles are then walked in order because run boundaries carry across them
		for (int base0 = b; base0 < e; base0 += 128) {
			ulonglong2 tile[4];
#pragma unroll
			for (int k = 0; k < 4; ++k) {
				const int i = base0 + 32 * k + lane;
				tile[k] = i < e ? __ldcs(&w.a[i]) : make_ulonglong2(0ull, 0ull);
			}
#pragma unroll
			for (int k = 0; k < 4; ++k) {
				const int base = base0 + 32 * k;
				if (base >= e) break;
				const int i = base + lane;
				const bool valid = i < e;
				const ulonglong2 ai = tile[k];
				unsigned long long xp = __shfl_up_sync(FULL, ai.x, 1);
				if (lane == 0) xp = prev_last_x;
				const bool start = valid && (i == b || ai.x - xp > mdx);
				prev_last_x = __shfl_sync(FULL, ai.x, 31);
				if (valid) {
					const int span = (int)((ai.y >> 32) & 0xff);
					sum += (unsigned)span;
					if ((ai.y >> 48) & 0xff) multiseg = 1;
				}
				const unsigned smask = __ballot_sync(FULL, start);
				if (valid) {
					// A run start gets f = span, p = -1, v = span (the fill kernels write all other anchors of a run).  A singleton
					// (a start whose successor starts the next run, or the last anchor of the read) needs no record at all: its
					// span travels in the mark word, t[i] = -2 - span (decoded in k_compact and k_export_fp), which saves the 20 scattered
					// bytes here and the 20 bytes the compaction would read back.  Lane 31 cannot see its successor and writes
					// the full record instead (equally valid).
					const int span = (int)((ai.y >> 32) & 0xff);
					const bool single = start && (i + 1 >= e || (lane < 31 && ((smask >> (lane + 1)) & 1u)));
					if (start && !single) { w.wrec[i] = make_int4((int)(unsigned)ai.x, (int)(unsigned)ai.y, span, -1); w.v[i] = span; }
					w.t[i] = single ? -2 - span : -1;
				}
				const unsigned below = smask & lanemask_lt(lane);
				const int prev_start = below ? base + (31 - __clz(below)) : carry_start;
				const bool emit = start && i != b && (i - prev_start) >= 2;   // this start closes the run [prev_start, i)
				if (emit) emit_unit(w, prev_start, i - prev_start, r);
				if (smask) carry_start = base + (31 - __clz(smask));
			}
		}
		for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(FULL, sum, d);
		multiseg = __any_sync(FULL, multiseg);
		if (lane == 0) {
			if (e - carry_start >= 2) emit_unit(w, carry_start, e - carry_start, r);
			// (float)sum_qspan / n  (chain.c:47): u64->float, int64->float, IEEE float divide
			w.avg[r] = e > b ? __fdiv_rn(__ull2float_rn(sum), __ll2float_rn((long long)(e - b))) : 0.0f;
			if (multiseg) atomicOr(&w.counters[CNT_MULTISEG], 1);
		}
	}
}

// ---------------------------------------------------------------------------------------------------------
// k_fill: the DP fill (chain.c:52-90), one warp per run, runs pulled from a device-side queue.
//
// For anchor i the 32 lanes score 32 predecessors j = jb - lane (descending j, chain.c:59) in parallel
// (SURVEY B.2), publish the marks t[p[j]] = i, synchronise, read t[j] (B.3), and then resolve the sequential
// state machine of chain.c:80-87 exactly (B.4): IMPROVE lanes are the strict prefix maxima, n_skip is replayed
// over the ballot masks segment by segment between IMPROVE lanes, and the break lane is the first MARKED-MISS
// lane that pushes n_skip past max_skip.  The winner is the last IMPROVE lane before the break, which is also
// the reference's "largest j wins ties" rule.
//
// Window test: within a run consecutive gaps are <= max_dist_x < 2^31, so the true distance of the first lane
// that falls out of the window is < 2^32 and the low 32 bits of x decide it exactly; lanes past it are dropped.
// ---------------------------------------------------------------------------------------------------------
static __device__ __forceinline__ int lin_cost(int dd, double avg)
{	// (int)(dd * .01 * avg_qspan): ((double)dd * 0.01) * (double)avg, truncation toward zero
	return __double2int_rz(__dmul_rn(__dmul_rn(__int2double_rn(dd), 0.01), avg));
}

// The sequential state machine of chain.c:80-87 replayed over one chunk's ballot masks.
//   imp : IMPROVE lanes (strict prefix maxima), mm : MARKED-MISS lanes (never overlap imp)
// Walks the segments between IMPROVE lanes; returns the break lane (or -1) and updates n_skip.  Warp-uniform.
static __device__ __forceinline__ int replay_skip(unsigned imp, unsigned mm, int max_skip, int lane, int &n_skip)
{
	unsigned rem = imp;
	int pos = 0;
	for (;;) {
		const int nxt = rem ? __ffs(rem) - 1 : 32;
		const unsigned seg = mm & mask_below(nxt) & ~mask_below(pos);
		const int cnt = __popc(seg);
		const int need = max(max_skip - n_skip + 1, 1);
		if (cnt >= need) {   // the need-th MARKED-MISS of this segment breaks (chain.c:84-85)
			const int c = __popc(seg & (lanemask_lt(lane) | (1u << lane)));
			const unsigned hit = __ballot_sync(FULL, ((seg >> lane) & 1u) && c == need);
			return __ffs(hit) - 1;
		}
		n_skip += cnt;
		if (nxt == 32) return -1;
		if (n_skip > 0) --n_skip;   // chain.c:82
		rem &= rem - 1;
		pos = nxt + 1;
	}
}

// ---- generic variant: cDNA / multi-segment scoring (chain.c:71-77); everything through global memory --------
__global__ void __launch_bounds__(128) k_fill_generic(Workspace w, mm2b200_params_t prm)
{
	// both variants are launched; the device-side flag decides which one does the work
	const bool need_generic = prm.is_cdna || (w.counters[CNT_MULTISEG] != 0);
	if (!need_generic) return;

	const int lane = threadIdx.x & 31;
	const int n_long = w.counters[CNT_LONG_UNITS];
	const int n_units = w.counters[CNT_UNITS] + n_long;
	const unsigned mdx = (unsigned)prm.max_dist_x;
	const int max_skip = prm.max_skip, bw = prm.bw;
	const int max_iter = prm.max_iter > 0 ? prm.max_iter : 0x7fffffff;
	const ulonglong2 *__restrict__ a = w.a;
	int4 *wrec = w.wrec;
	int32_t *v = w.v, *t = w.t;

	for (;;) {
		int u = 0;
		if (lane == 0) u = atomicAdd(&w.counters[CNT_QUEUE], 1);
		u = __shfl_sync(FULL, u, 0);
		if (u >= n_units) break;
		const Unit un = fetch_unit(w, u, n_long);
		const int s = un.start, e = un.start + un.len;
		const double avg = (double)w.avg[un.read];
		long long n_visited = 0;   // loop bodies the reference executes (chain.c:59-88), the metric's numerator
		long long n_chunks = 0;    // 32-lane chunks evaluated (lane efficiency = visited / (32 * chunks))

		for (int i = s + 1; i < e; ++i) {
			const ulonglong2 ai = a[i];
			const unsigned xi = (unsigned)ai.x;
			const int qi = (int)(unsigned)ai.y;
			const int span = (int)((ai.y >> 32) & 0xff);
			const int sidi = (int)((ai.y >> 48) & 0xff);
			int max_f = span, max_j = -1, n_skip = 0;
			bool done = false;
			const int jlo = (i - s > max_iter) ? i - max_iter : s;      // optional max_iter clamp (0 = off)

			for (int jb = i - 1; jb >= jlo && !done; jb -= 32) {
				const int j = jb - lane;
				const bool valid = j >= jlo;
				unsigned long long ajy = 0;
				int4 rj = make_int4(0, 0, 0, -1);
				if (valid) { ajy = a[j].y; rj = wrec[j]; }
				const int fj = rj.z, pj = rj.w;
				const unsigned dr = xi - (unsigned)rj.x;
				const unsigned wmask = __ballot_sync(FULL, valid && dr <= mdx);
				int nwin = 32;
				if (wmask != FULL) { nwin = __ffs(~wmask) - 1; done = true; }
				if (nwin == 0) break;
				n_visited += nwin; ++n_chunks;
				const bool act = lane < nwin;

				// ---- pair score (chain.c:60-79) ----
				const int dq = qi - rj.y;
				const bool same = sidi == (int)((ajy >> 48) & 0xff);
				bool skip = !act || (same && dr == 0) || dq <= 0;
				skip = skip || (same && dq > prm.max_dist_y) || dq > prm.max_dist_x;
				const int dd = abs((int)dr - dq);
				skip = skip || (same && dd > bw);
				skip = skip || (prm.n_segs > 1 && !prm.is_cdna && same && dr > (unsigned)prm.max_dist_y);
				int sc = min(min(dq, (int)dr), span);
				const int log_dd = dd ? 31 - __clz(dd) : 0;
				const int c_lin = lin_cost(dd, avg);
				if (prm.is_cdna || !same) {
					if (!same && dr == 0) ++sc;
					else if ((int)dr > dq || !same) sc -= min(c_lin, log_dd);
					else sc -= c_lin + (log_dd >> 1);
				} else sc -= c_lin + (log_dd >> 1);
				sc += fj;

				// ---- marks (chain.c:83,87): publish, synchronise, read ----
				if (!skip && pj >= 0) t[pj] = i;
				__syncwarp();
				const bool marked = !skip && t[j] == i;

				// ---- IMPROVE lanes = strict prefix maxima over the visit order (chain.c:80) ----
				int pm = skip ? (int)0x80000000 : sc;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const int o = __shfl_up_sync(FULL, pm, d);
					if (lane >= d) pm = max(pm, o);
				}
				int ex = __shfl_up_sync(FULL, pm, 1);
				ex = lane == 0 ? max_f : max(ex, max_f);
				const bool improve = !skip && sc > ex;
				const unsigned imp = __ballot_sync(FULL, improve);
				const unsigned mm = __ballot_sync(FULL, marked && !improve);

				const int brk = replay_skip(imp, mm, max_skip, lane, n_skip);
				const unsigned impb = brk >= 0 ? (imp & mask_below(brk)) : imp;
				if (impb) {
					const int L = 31 - __clz(impb);
					max_f = __shfl_sync(FULL, sc, L);
					max_j = jb - L;
				}
				if (brk >= 0) { done = true; n_visited -= nwin - (brk + 1); }
			}
			if (lane == 0) {
				int vi = max_f;
				if (max_j >= 0) { const int vj = v[max_j]; if (vj > max_f) vi = vj; }   // chain.c:90
				wrec[i] = make_int4((int)xi, qi, max_f, max_j);
				v[i] = vi;
			}
			__syncwarp();
		}
		if (lane == 0) { atomicAdd(w.visited, (unsigned long long)n_visited); atomicAdd(w.visited + 1, (unsigned long long)n_chunks); }
	}
}

// ---- fast variant: single segment, not cDNA (every long-read preset) ------------------------------------------
//
// Issue-slot budget is what bounds this kernel (ncu: 3.0 of 4 issue slots/clk busy, HBM < 5 %), so the design
// minimises instructions per (anchor, 32-predecessor chunk):
//  * each warp keeps the last 64 anchors of its run in a shared-memory ring of packed records {x, q, f, p} (+ v):
//    one LDS.128 feeds a lane; chunks deeper than 64 read the same records from global memory (L1/L2);
//  * the gap cost (int)(dd*.01*avg_qspan) + (ilog2(dd)>>1) of every dd <= bw (chain.c:66 guarantees dd <= bw for
//    a scored pair) comes from a per-warp shared-memory table built with the reference's double arithmetic and
//    rebuilt only when avg_qspan changes (TABLE; bw <= 1023);
//  * marks inside the chunk never touch memory: lane k' contributes bit (jb - p[j']) to a warp OR-reduce
//    (redux.sync); only marks that land beyond the chunk go through t[] in global memory;
//  * IMPROVE lanes: one redux.max finds the chunk maximum M and the first lane L that attains it; if no earlier
//    lane beats the running maximum, L is the only IMPROVE lane (the common case) and no prefix scan is needed.
#define FILL_WARPS 4
// Dynamic shared memory of k_fill_fast<RING>, byte offsets from the CTA's (RING*16-byte aligned) base.  Per warp:
//   ring       RING x int4 {x, q, f, p} of the last RING anchors
//   companion  RING x 16 B, same slot addressing: word 0 = v[] (chain.c:90), word 1 = mark tag (t[] of chain.c:83,87 for
//              the anchors that are still in the ring; marks on older anchors go through t[] in global memory)
//   abuf       32 x int4 {x, q, span, -} of the current 32-aligned block of anchors
//   table      16-bit gap costs, tab_entries per warp
// RING = 64 (two chunks; 3.5 KB per warp, 10 CTAs per SM) for ordinary reads, RING = 256 (eight chunks; 9.5 KB per warp)
// for sub-batches of long reads, whose windows hold hundreds of predecessors (map-pb: 157 visited pairs per anchor).
template <int RING> struct FillSmem {
	static constexpr int ring = 0;
	static constexpr int comp = FILL_WARPS * RING * 16;
	static constexpr int abuf = 2 * FILL_WARPS * RING * 16;
	static constexpr int tab = abuf + FILL_WARPS * 32 * 16;
	static constexpr unsigned mask = RING * 16u - 1u;
};

// explicit 32-bit shared addresses: every access is one LDS/STS with an immediate offset, no generic-address math
template <int OFF> static __device__ __forceinline__ int4 lds128(unsigned addr)
{
	int4 v;
	asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4+%5];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr), "n"(OFF));
	return v;
}
template <int OFF> static __device__ __forceinline__ int lds32(unsigned addr)
{
	int v;
	asm volatile("ld.shared.s32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
	return v;
}
static __device__ __forceinline__ int lds_u16(unsigned addr)
{
	unsigned short v;
	asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
	return (int)v;
}
template <int OFF> static __device__ __forceinline__ void sts128(unsigned addr, int4 v)
{
	asm volatile("st.shared.v4.s32 [%0+%1], {%2,%3,%4,%5};" :: "r"(addr), "n"(OFF), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int OFF> static __device__ __forceinline__ void sts32(unsigned addr, int v)
{
	asm volatile("st.shared.s32 [%0+%1], %2;" :: "r"(addr), "n"(OFF), "r"(v) : "memory");
}
static __device__ __forceinline__ void sts_u16(unsigned addr, int v)
{
	asm volatile("st.shared.u16 [%0], %1;" :: "r"(addr), "h"((unsigned short)v) : "memory");
}

// a + IMM issued on the FMA pipe (IMAD a, one, IMM) instead of the ALU pipe (VIADD): in the fill kernel the ALU pipe
// (compares, logic, selects, min/max: one warp instruction every two cycles) is the busier of the two integer pipes
// (ncu capture k: ALU 80 %, FMA 20 %), so plain add-immediates are moved over.  `one` is a register holding 1 that is
// opaque to the optimiser.
template <int IMM> static __device__ __forceinline__ int addi_fma(int a, int one)
{
	int d;
	asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "n"(IMM));
	return d;
}
// 2 * a + b as one IMAD (table address)
static __device__ __forceinline__ unsigned mad2_fma(unsigned a, unsigned b)
{
	unsigned d;
	asm("mad.lo.u32 %0, %1, 2, %2;" : "=r"(d) : "r"(a), "r"(b));
	return d;
}
template <int IMM> static __device__ __forceinline__ unsigned addi_fma(unsigned a, int one)
{
	return (unsigned)addi_fma<IMM>((int)a, one);
}

// 1 << k with PTX shift semantics: 0 for k >= 32 (C++ leaves that undefined and the compiler would emit a wrapping shift)
static __device__ __forceinline__ unsigned bit_or_zero(unsigned k)
{
	unsigned r;
	asm("shl.b32 %0, 1, %1;" : "=r"(r) : "r"(k));
	return r;
}

// One 32-predecessor chunk of anchor i (lane = jb - j).  State carried across the chunks of one anchor:
//   max_f / max_j / vbest  running best predecessor (chain.c:80-81) and its peak score v[]
//   budget                 marked misses still allowed before the break: max_skip + 1 - n_skip (chain.c:82-85)
// Returns nonzero when the scan of anchor i is finished (break, or the window ended inside the chunk).
// `far` reports lanes whose mark lands beyond this chunk; the caller publishes those to t[] only if the scan goes on.
template <bool COUNT>
static __device__ __noinline__ int scan_chunk(int &max_f, int &max_j, int &vbest, int &budget, long long &visited, bool &far,
                                                  const int jb, const int jlo, const int lane,
                                                  const int4 rj, const int vj, const bool marked_mem,
                                                  const unsigned xi, const int qi, const int span,
                                                  const unsigned mdx, const unsigned dr_max, const unsigned dq_max, const int bw,
                                                  const int cost, const int B0, const int max_skip)
{
	const int j = jb - lane;
	const unsigned dr = xi - (unsigned)rj.x;
	bool inw = (j >= jlo) && (dr <= mdx);
	int done = 0, nwin = 32;
	far = false;
	if (!__all_sync(FULL, inw)) {                         // the window (chain.c:58) ends inside this chunk
		nwin = __ffs(~__ballot_sync(FULL, inw)) - 1;      // lanes past the first one out of the window are dropped
		done = 1;
		if (nwin == 0) return 1;
		inw = lane < nwin;
	}
	if (COUNT) visited += nwin;

	// ---- pair score (chain.c:60-79, same-segment non-cDNA branch) ----
	const int dq = qi - rj.y;
	const int dd = abs((int)dr - dq);
	// dr == 0 || dr > dr_max  <=>  dr-1 >= dr_max (unsigned);  dq <= 0 || dq > dq_max  <=>  dq-1 >= dq_max (unsigned)
	const bool skip = !inw || (dr - 1u) >= dr_max || ((unsigned)dq - 1u) >= dq_max || dd > bw;
	const int sc = min(min(dq, (int)dr), span) - cost + rj.z;
	const int scx = skip ? (int)0x80000000 : sc;

	// ---- marks (chain.c:83,87): inside the chunk via a warp OR-reduce, beyond it via t[] ----
	const int pj = rj.w;
	const unsigned kt = (unsigned)(jb - pj);              // lane that scores anchor p[j], if it is in this chunk
	const bool mk = !skip && pj >= 0;
	const unsigned mloc = __reduce_or_sync(FULL, (mk && kt < 32u) ? (1u << (kt & 31u)) : 0u);
	far = mk && kt >= 32u;
	const bool marked = ((mloc >> lane) & 1u) || marked_mem;

	// ---- IMPROVE lanes (chain.c:80): strict prefix maxima of the scores in visit order ----
	const int M = __reduce_max_sync(FULL, scx);
	unsigned imp = 0;
	if (M > max_f) {
		const unsigned eq = __ballot_sync(FULL, scx == M);
		const unsigned gt = __ballot_sync(FULL, scx > max_f);
		const unsigned low = eq & (0u - eq);              // first lane that attains the chunk maximum
		if ((gt & (low - 1u)) == 0u) imp = low;           // nobody before it beats the running maximum: the only IMPROVE
		else {                                            // several candidates: full exclusive prefix-max scan
			int pm = scx;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const int o = __shfl_up_sync(FULL, pm, d);
				if (lane >= d) pm = max(pm, o);
			}
			int ex = __shfl_up_sync(FULL, pm, 1);
			ex = lane == 0 ? max_f : max(ex, max_f);
			imp = __ballot_sync(FULL, scx > ex);
		}
	}
	const unsigned mm = __ballot_sync(FULL, marked && !skip) & ~imp;   // MARKED-MISS lanes

	// ---- replay n_skip and the break (chain.c:82-85) ----
	int brk = -1;
	if ((imp & (imp - 1u)) == 0u) {
		// at most one IMPROVE lane L: [marked misses before L] [L] [marked misses after L]; imp-1 = lanes below L
		const unsigned below = imp - 1u;                  // imp == 0 -> all lanes
		const unsigned seg1 = mm & below;
		const int c1 = __popc(seg1);
		if (c1 >= budget) {
			done = 1;
			if (COUNT) {
				const int c = __popc(seg1 & (lanemask_lt(lane) | (1u << lane)));
				brk = __ffs(__ballot_sync(FULL, ((seg1 >> lane) & 1u) && c == budget)) - 1;
			}
		} else {
			budget -= c1;
			if (imp) {
				const int L = __popc(below);
				max_f = M; max_j = jb - L;
				vbest = __shfl_sync(FULL, vj, L);
				budget = min(budget + 1, B0);             // if (n_skip > 0) --n_skip
				const unsigned seg2 = mm & ~(imp | below);
				const int c2 = __popc(seg2);
				if (c2 >= budget) {
					done = 1;
					if (COUNT) {
						const int c = __popc(seg2 & (lanemask_lt(lane) | (1u << lane)));
						brk = __ffs(__ballot_sync(FULL, ((seg2 >> lane) & 1u) && c == budget)) - 1;
					}
				} else budget -= c2;
			}
		}
	} else {
		int n_skip = B0 - budget;
		brk = replay_skip(imp, mm, max_skip, lane, n_skip);
		budget = B0 - n_skip;
		const unsigned impb = brk >= 0 ? (imp & mask_below(brk)) : imp;
		if (impb) {
			const int L = 31 - __clz(impb);
			max_f = __shfl_sync(FULL, sc, L);
			max_j = jb - L;
			vbest = __shfl_sync(FULL, vj, L);
		}
		if (brk >= 0) done = 1;
	}
	if (COUNT && brk >= 0) visited -= nwin - (brk + 1);
	return done;
}

// The common case of a chunk as straight-line code: at most one IMPROVE lane and a window that is exact per lane
// (max_dist_x < 2^26, so the 33 gaps down to a chunk's last lane can never wrap the low 32 bits of x).  Everything is computed with selects; the only
// branch leaves for the general routine above when several lanes before the chunk maximum beat the running maximum.
template <bool COUNT, bool WIDE>
static __device__ __forceinline__ int scan_chunk_fast(int &max_f, int &max_j, int &vbest, int &budget, long long &visited, bool &far,
                                                       const int jb, const int jlo, const int lane,
                                                       const int4 rj, const int vj, const bool marked_mem,
                                                       const unsigned xi, const int qi, const int span,
                                                       const unsigned mdx, const unsigned dr_max, const unsigned dq_max, const int bw,
                                                       const int dd, const int cost, const int B0, const int max_skip, const int one)
{
	const int j = jb - lane;
	const unsigned dr = xi - (unsigned)rj.x;
	const int dq = qi - rj.y;
	const bool inw = (j >= jlo) && (dr <= mdx);
	const bool skip = !inw || addi_fma<-1>(dr, one) >= dr_max || addi_fma<-1>((unsigned)dq, one) >= dq_max || dd > bw;
	const int sc = min(min(dq, (int)dr), span) - cost + rj.z;
	const int scx = skip ? (int)0x80000000 : sc;
	const int pj = rj.w;
	const unsigned kt = (unsigned)(jb - pj);
	const bool mk = !skip && pj >= 0;
	const unsigned mloc = __reduce_or_sync(FULL, mk ? bit_or_zero(kt) : 0u);
	const int M = __reduce_max_sync(FULL, scx);
	const bool all_in = __all_sync(FULL, inw);
	const unsigned eq = __ballot_sync(FULL, scx == M);
	const unsigned gt = __ballot_sync(FULL, scx > max_f);
	const unsigned low = eq & (0u - eq);
	if (WIDE)                               // window too wide for the per-lane test: general routine (prefix of in-window lanes)
		return scan_chunk<COUNT>(max_f, max_j, vbest, budget, visited, far, jb, jlo, lane, rj, vj, marked_mem,
		                         xi, qi, span, mdx, dr_max, dq_max, bw, cost, B0, max_skip);
	far = mk && kt >= 32u;
	const bool marked = ((mloc >> lane) & 1u) || marked_mem;
	if ((gt & addi_fma<-1>(low, one)) != 0u) {
		// several lanes before the chunk maximum beat the running maximum => at least two IMPROVE lanes:
		// exclusive prefix-max scan for the exact set, then the general segment-by-segment replay
		int pm = scx;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const int o = __shfl_up_sync(FULL, pm, d);
			if (lane >= d) pm = max(pm, o);
		}
		int ex = __shfl_up_sync(FULL, pm, 1);
		ex = lane == 0 ? max_f : max(ex, max_f);
		const unsigned impm = __ballot_sync(FULL, scx > ex);
		const unsigned mmm = __ballot_sync(FULL, marked && !skip) & ~impm;
		int n_skip = B0 - budget;
		const int brk = replay_skip(impm, mmm, max_skip, lane, n_skip);
		budget = B0 - n_skip;
		const unsigned impb = brk >= 0 ? (impm & mask_below(brk)) : impm;
		if (impb) {
			const int Lb = 31 - __clz(impb);
			max_f = __shfl_sync(FULL, sc, Lb);
			max_j = jb - Lb;
			vbest = __shfl_sync(FULL, vj, Lb);
		}
		if (COUNT) {
			const int nwin = __popc(__ballot_sync(FULL, inw));
			visited += nwin;
			if (brk >= 0) visited -= nwin - (brk + 1);
		}
		return (brk >= 0 || !all_in) ? 1 : 0;
	}
	const unsigned imp = M > max_f ? low : 0u;            // the only IMPROVE lane, if any
	const unsigned below = addi_fma<-1>(imp, one);        // lanes visited before it (all lanes when there is none)
	const unsigned mm = __ballot_sync(FULL, marked && !skip) & ~imp;
	const unsigned seg1 = mm & below, seg2 = mm & ~(imp | below);
	const int c1 = __popc(seg1), c2 = __popc(seg2);
	const bool brk1 = c1 >= budget;                       // break before the IMPROVE lane (chain.c:84-85)
	const bool take = imp != 0u && !brk1;
	const int L = __popc(below) & 31;
	const int vL = __shfl_sync(FULL, vj, L);
	const int b1 = budget - c1, b2 = min(b1 + 1, B0);     // chain.c:82: if (n_skip > 0) --n_skip
	const bool brk2 = take && c2 >= b2;
	if (COUNT) {
		const int nwin = __popc(__ballot_sync(FULL, inw));
		visited += nwin;
		if (brk1 || brk2) {
			const unsigned seg = brk1 ? seg1 : seg2;
			const int need = brk1 ? budget : b2;
			const int c = __popc(seg & (lanemask_lt(lane) | (1u << lane)));
			const int brk = __ffs(__ballot_sync(FULL, ((seg >> lane) & 1u) && c == need)) - 1;
			visited -= nwin - (brk + 1);
		}
	}
	max_f = take ? M : max_f;
	max_j = take ? jb - L : max_j;
	vbest = take ? vL : vbest;
	budget = take ? b2 - c2 : b1;
	return (brk1 || brk2 || !all_in) ? 1 : 0;
}

// Chunks beyond the first are mostly dead: on the BASELINE workloads 46-57 % of them have every lane SKIPped
// (chain.c:63-67) and 93-97 % have at most two scored lanes -- they belong to off-chain anchors that walk through a
// long window of predecessors on other diagonals.  Such a chunk needs no warp-wide replay: an all-SKIP chunk changes
// nothing, and a few scored lanes are stepped through the reference's state machine one by one, in visit order, with
// warp-uniform scalar code (two shuffles per scored lane).  Returns 0 = scan goes on, 1 = anchor finished,
// 2 = too many scored lanes, nothing was changed: run the dense routine on this chunk.
#ifndef DEEP_BATCH
#define DEEP_BATCH 4      // chunks of a deep window requested together (k_fill_fast<.., DEEP = true>)
#endif
#define SPARSE_MAX 4
// a mark on anchor pj while anchor i is scanned (chain.c:87): in the ring's tag word while pj is still in the ring
template <int RING>
static __device__ __forceinline__ void publish_mark(const int pj, const int i, const unsigned ring_w, int32_t *t)
{
	if (i - pj <= RING) sts32<FillSmem<RING>::comp + 4>(ring_w + (((unsigned)pj & (RING - 1)) << 4), i);
	else t[pj] = i;
}

template <bool TABLE, bool COUNT, bool WIDE, int RING>
static __device__ __forceinline__ int scan_chunk_sparse(int &max_f, int &max_j, int &vbest, int &budget, long long &visited,
                                                         const int i, const int jb, const int jlo, const int lane,
                                                         const int4 rj, const bool marked_mem,
                                                         const unsigned xi, const int qi, const int span,
                                                         const unsigned mdx, const unsigned dr_max, const unsigned dq_max, const int bw,
                                                         const int B0, const unsigned tab_w, const double avg,
                                                         int32_t *t, const int32_t *vg, const unsigned ring_w, bool &pub)
{
	const int j = jb - lane;
	const unsigned dr = xi - (unsigned)rj.x;
	const int dq = qi - rj.y;
	const int dd = abs((int)dr - dq);
	bool inw = (j >= jlo) && (dr <= mdx);
	int done, nwin = 32;
	if (WIDE || COUNT) {
		const unsigned wb = __ballot_sync(FULL, inw);
		done = wb != FULL;
		if (done) {                                       // the window (chain.c:58) ends inside this chunk: lanes past the
			nwin = __ffs(~wb) - 1;                        // first one outside are dropped (exact for wide windows too)
			inw = lane < nwin;
		}
	} else done = !__all_sync(FULL, inw);                 // max_dist_x < 2^26: the per-lane test is exact
	const bool skip = !inw || (dr - 1u) >= dr_max || ((unsigned)dq - 1u) >= dq_max || dd > bw;
	unsigned live = __ballot_sync(FULL, !skip);
	if (COUNT) visited += nwin;
	if (live == 0u) return done;                          // nothing scored: no IMPROVE, no MISS, no mark
	if (__popc(live) > SPARSE_MAX) { if (COUNT) visited -= nwin; return 2; }

	int sc = 0;
	if (!skip) {
		const int cost = TABLE ? lds_u16(tab_w + (unsigned)dd * 2u) : lin_cost(dd, avg) + ((31 - __clz(dd | 1)) >> 1);
		sc = min(min(dq, (int)dr), span) - cost + rj.z;   // chain.c:68-79
	}
	const unsigned mmem = __ballot_sync(FULL, marked_mem);
	unsigned lm = 0;                                      // marks set by earlier scored lanes of this chunk
	do {
		const int L = __ffs(live) - 1;
		live &= live - 1u;
		const int scL = __shfl_sync(FULL, sc, L), pjL = __shfl_sync(FULL, rj.w, L);
		if (scL > max_f) {                                // chain.c:80-82
			max_f = scL; max_j = jb - L;
			vbest = (i - max_j <= RING) ? lds32<FillSmem<RING>::comp>(ring_w + (((unsigned)max_j & (RING - 1)) << 4)) : vg[max_j];
			budget = min(budget + 1, B0);
		} else if (((mmem | lm) >> L) & 1u) {             // chain.c:83-85
			if (--budget <= 0) {
				if (COUNT) visited -= nwin - (L + 1);
				return 1;
			}
		}
		if (pjL >= 0) {                                   // chain.c:87
			const unsigned kt = (unsigned)(jb - pjL);
			if (kt < 32u) lm |= 1u << kt;
			else { pub = true; if (lane == 0) publish_mark<RING>(pjL, i, ring_w, t); }    // pub: warp-uniform "a mark left this chunk"
		}
	} while (live);
	return done;
}

// FULL = instrumented / extended variant: counts visited pairs and honours max_iter; the plain variant does neither
// (register caps were tried: 40 registers / 12 CTAs per SM 5.70 ms, 32 / 16 CTAs 6.09 ms, 67 / 7 CTAs 5.07 ms, default 48 / 10 CTAs 4.89 ms)
template <bool TABLE, bool FULLV, bool WIDE, int RING, bool DEEP>
__global__ void __launch_bounds__(FILL_WARPS * 32) k_fill_fast(Workspace w, mm2b200_params_t prm, int tab_entries)
{
	extern __shared__ int4 smem4[];
	typedef FillSmem<RING> SM;
	if (prm.is_cdna || (w.counters[CNT_MULTISEG] != 0)) return;   // the generic kernel handles those

	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	// base aligned to the ring size: a warp's ring is then a naturally aligned window and ring addresses wrap with one LOP3
	const unsigned sbase = ((unsigned)__cvta_generic_to_shared(smem4) + SM::mask) & ~SM::mask;
	unsigned ring_w = sbase + SM::ring + wib * (RING * 16);
	unsigned abuf_w = sbase + SM::abuf + wib * (32 * 16);
	unsigned tab_w = sbase + SM::tab + wib * tab_entries * 2;
	// opaque to the optimiser: keep the three per-warp base addresses in registers instead of re-deriving them per access
	asm volatile("" : "+r"(ring_w), "+r"(abuf_w), "+r"(tab_w));
	const int one = 1 + (int)threadIdx.y;         // see addi_fma (threadIdx.y is 0, which ptxas cannot know)

	const int n_long = w.counters[CNT_LONG_UNITS];
	const int n_units = w.counters[CNT_UNITS] + n_long;
	unsigned mdx = (unsigned)prm.max_dist_x;
	const unsigned dq_max = (unsigned)(prm.max_dist_y < prm.max_dist_x ? prm.max_dist_y : prm.max_dist_x);   // chain.c:64
	const unsigned dr_max = prm.n_segs > 1 ? (unsigned)prm.max_dist_y : 0xffffffffu;                             // chain.c:67
	const int max_skip = prm.max_skip;
	int bw = prm.bw;
	// per-thread registers instead of a constant-bank re-read per anchor (threadIdx.y is 0; the compiler cannot fold it)
	mdx += threadIdx.y; bw += (int)threadIdx.y;
	const int B0 = max(max_skip + 1, 1);
	const int max_iter = (FULLV && prm.max_iter > 0) ? prm.max_iter : 0x7fffffff;
	const ulonglong2 *__restrict__ a = w.a;
	int4 *wrec = w.wrec;
	int32_t *vg = w.v, *t = w.t;
	unsigned tab_key = 0;
	bool have_tab = false;

	// no mark tag may match an anchor index before it is written (tags are batch-wide anchor indices, all >= 0)
	for (int k = lane; k < RING; k += 32) sts32<SM::comp + 4>(ring_w + (unsigned)k * 16u, -1);
	__syncwarp();

	for (;;) {
		// next run from the device-side queue; the redux makes the index provably warp-uniform for the compiler
		int u = 0;
		if (lane == 0) u = atomicAdd(&w.counters[CNT_QUEUE], 1);
		u = __reduce_max_sync(FULL, u);
		if (u >= n_units) break;
		const Unit un = fetch_unit(w, u, n_long);
		const int s = un.start, e = un.start + un.len;
		const float avgf = w.avg[un.read];
		const double avg = (double)avgf;
		long long visited = 0, n_chunks = 0;

		if (TABLE && (!have_tab || __float_as_uint(avgf) != tab_key)) {
			__syncwarp();
			for (int dd = lane; dd <= bw; dd += 32)
				sts_u16(tab_w + dd * 2, lin_cost(dd, avg) + ((31 - __clz(dd | 1)) >> 1));
			tab_key = __float_as_uint(avgf);
			have_tab = true;
		}
		// seed the ring with the run's first anchor (f = span, p = -1, v = span: chain.c:56,89-90 with no predecessor)
		{
			const ulonglong2 a0 = a[s];
			const int sp = (int)((a0.y >> 32) & 0xff);
			const unsigned slot = ring_w + (unsigned)(s & (RING - 1)) * 16u;
			sts128<0>(slot, make_int4((int)(unsigned)a0.x, (int)(unsigned)a0.y, sp, -1));
			sts32<SM::comp>(slot, sp);
		}

		// anchors are processed in 32-aligned blocks: lane l loads a[blk + l] once, every step broadcasts its own from
		// shared memory, and the block's results go back to global memory with one coalesced store per lane
		for (int blk = (s + 1) & ~31; blk < e; blk += 32) {
			__syncwarp();
			if (blk + lane < e) {
				const ulonglong2 ab = a[blk + lane];
				sts128<0>(abuf_w + lane * 16, make_int4((int)(unsigned)ab.x, (int)(unsigned)ab.y, (int)((ab.y >> 32) & 0xff), 0));
			}
			__syncwarp();
			const int i1 = min(blk + 32, e);
			const int i0 = max(blk, s + 1);
			// running shared addresses: anchor i in the block buffer, this lane's chunk-0 record (anchor i-1-lane), slot of i
			unsigned aaddr = abuf_w + (unsigned)(i0 & 31) * 16u;
			unsigned raddr = ring_w + (unsigned)((i0 - 1 - lane) & (RING - 1)) * 16u;
			unsigned waddr = ring_w + (unsigned)(i0 & (RING - 1)) * 16u;
		for (int i = i0; i < i1; ++i) {
			const int4 ai = lds128<0>(aaddr);
			const unsigned xi = (unsigned)ai.x;
			const int qi = ai.y, span = ai.z;
			const int jlo = FULLV ? max(s, i - max_iter) : s;
			int max_f = span, max_j = -1, vbest = (int)0x80000000, budget = B0;
			bool far;

			// chunk 0: anchors i-1 .. i-32 from the ring; no earlier chunk can have marked them
			int4 rj = lds128<0>(raddr);
			int vj = lds32<SM::comp>(raddr);
			int dd0 = abs((int)(xi - (unsigned)rj.x) - (qi - rj.y));
			int cost = TABLE ? lds_u16(mad2_fma(min((unsigned)dd0, (unsigned)bw), tab_w)) : lin_cost(dd0, avg) + ((31 - __clz(dd0 | 1)) >> 1);
			if (FULLV) ++n_chunks;
			int done = scan_chunk_fast<FULLV, WIDE>(max_f, max_j, vbest, budget, visited, far, addi_fma<-1>(i, one), jlo, lane, rj, vj, false,
			                                         xi, qi, span, mdx, dr_max, dq_max, bw, dd0, cost, B0, max_skip, one);
			if (!done && addi_fma<-33>(i, one) >= jlo) {
				// marks of chunk 0 that fell beyond it: tag word of the ring slot (or t[] for anchors that left the ring)
				if (__any_sync(FULL, far)) { if (far) publish_mark<RING>(rj.w, i, ring_w, t); }
				// one chunk beyond the first: records and marks from the ring (chunks 1 .. RING/32-1) or from global memory
				// (the records flushed there in 32-anchor blocks; L1/L2 hits)
				// one chunk beyond the first, records rk and memory marks mg already loaded; pub reports marks that left the chunk
				auto process = [&](const bool in_ring, const unsigned ra, const int jb, const int jc, const int4 rk, const bool mg, bool &pub) -> int {
					if (FULLV) ++n_chunks;
					int dn = scan_chunk_sparse<TABLE, FULLV, WIDE, RING>(max_f, max_j, vbest, budget, visited, i, jb, jlo, lane, rk, mg,
					                                       xi, qi, span, mdx, dr_max, dq_max, bw, B0, tab_w, avg, t, vg, ring_w, pub);
					if (dn == 2) {                                    // many scored lanes: warp-wide replay
						const int vk = in_ring ? lds32<SM::comp>(ra) : vg[jc];
						const int ddk = abs((int)(xi - (unsigned)rk.x) - (qi - rk.y));
						const int ck = TABLE ? lds_u16(tab_w + min((unsigned)ddk, (unsigned)bw) * 2u) : lin_cost(ddk, avg) + ((31 - __clz(ddk | 1)) >> 1);
						bool fk;
						dn = scan_chunk_fast<FULLV, WIDE>(max_f, max_j, vbest, budget, visited, fk, jb, jlo, lane, rk, vk, mg,
						                                   xi, qi, span, mdx, dr_max, dq_max, bw, ddk, ck, B0, max_skip, one);
						if (!dn) {                                    // (the scalar path publishes its own)
							if (__any_sync(FULL, fk)) pub = true;
							if (fk) publish_mark<RING>(rk.w, i, ring_w, t);
						}
					}
					return dn;
				};
				auto deeper = [&](const bool in_ring, const unsigned ra, const int jb) -> int {
					__syncwarp();                                     // marks published by the previous chunk are visible
					const int jc = max(jb - lane, jlo);               // clamped: lanes below jlo are masked (j >= jlo in the scan routines)
					int4 rk; bool mg;
					if (in_ring) { rk = lds128<0>(ra); mg = lds32<SM::comp + 4>(ra) == i; }
					else { rk = wrec[jc]; mg = t[jc] == i; }
					bool pub = false;
					return process(in_ring, ra, jb, jc, rk, mg, pub);
				};
				unsigned ra = ring_w | (addi_fma<-512>(raddr, one) & SM::mask);
				done = deeper(true, ra, i - 33);                      // chunk 1 is always in the ring
				int jb = i - 65;
				if (RING > 64) {
#pragma unroll 1
					for (int c = 2; c < RING / 32 && !done && jb >= jlo; ++c, jb -= 32) {
						ra = ring_w | (addi_fma<-512>(ra, one) & SM::mask);
						done = deeper(true, ra, jb);
					}
				}
				if (DEEP) {
					// Long reads: most of the scan is out here, one L2 / HBM round trip per chunk (ncu, map-pb: 57 % of the stall
					// samples on these two loads, L2 hit rate 49 %).  While at least DEEP_BATCH whole chunks of the window remain, their
					// records and marks are requested together and worked off in order; a mark that one of them publishes may be
					// missing from the early load of the later ones, so those reload their marks (rare: the deep part of a window
					// is mostly off the diagonal).  An early exit wastes at most DEEP_BATCH - 1 chunk loads.
#pragma unroll 1
					while (!done && jb >= jlo) {
						if (jb - (32 * DEEP_BATCH - 1) >= jlo) {
							__syncwarp();
							int4 rk[DEEP_BATCH]; int tk[DEEP_BATCH];
#pragma unroll
							for (int c = 0; c < DEEP_BATCH; ++c) { rk[c] = wrec[jb - 32 * c - lane]; tk[c] = t[jb - 32 * c - lane]; }
							bool pub = false;
#pragma unroll
							for (int c = 0; c < DEEP_BATCH; ++c) {
								if (!done) {
									const int jc = jb - 32 * c - lane;
									if (c) { __syncwarp(); if (pub) tk[c] = t[jc]; }
									done = process(false, 0u, jb - 32 * c, jc, rk[c], tk[c] == i, pub);
								}
							}
							jb -= 32 * DEEP_BATCH;
						} else {
							done = deeper(false, 0u, jb);
							jb -= 32;
						}
					}
				} else {
#pragma unroll 1
					for (; !done && jb >= jlo; jb -= 32) done = deeper(false, 0u, jb);
				}
			}
			// chain.c:89-90: every lane holds the same result; all write it to the ring (same address, same data)
			sts128<0>(waddr, make_int4((int)xi, qi, max_f, max_j));
			sts32<SM::comp>(waddr, max(vbest, max_f));
			aaddr = addi_fma<16>(aaddr, one);
			raddr = ring_w | (addi_fma<16>(raddr, one) & SM::mask);
			waddr = ring_w | (addi_fma<16>(waddr, one) & SM::mask);
			// (no warp sync needed: every lane stored the same words, so a later load by any lane sees at least its own store)
		}
			// flush the block (deep windows and the later kernels read the records from global memory)
			{
				const int idx = blk + lane;
				if (idx > s && idx < i1) {
					const unsigned slot = ring_w + (unsigned)(idx & (RING - 1)) * 16u;
					wrec[idx] = lds128<0>(slot);
					vg[idx] = lds32<SM::comp>(slot);
				}
			}
		}
		if (FULLV && lane == 0) { atomicAdd(w.visited, (unsigned long long)visited); atomicAdd(w.visited + 1, (unsigned long long)n_chunks); }
	}
}

// ---------------------------------------------------------------------------------------------------------
// compaction (chain.c:93-122)
// ---------------------------------------------------------------------------------------------------------
static __device__ __forceinline__ void store_rec(new_seed *dst, ulonglong2 seed, int32_t p, int32_t f)
{
	unsigned long long *q = reinterpret_cast<unsigned long long*>(dst);
	q[0] = seed.x; q[1] = seed.y;
	q[2] = (unsigned long long)(unsigned)p | ((unsigned long long)(unsigned)f << 32);
}

__global__ void __launch_bounds__(256) k_compact(Workspace w, int32_t min_sc)
{
	const int lane = threadIdx.x & 31;
	const int gw = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
	const int nw = (int)((gridDim.x * (unsigned)blockDim.x) >> 5);
	for (int r = gw; r < w.n_reads; r += nw) {
		const int b = (int)w.offs[r], e = (int)w.offs[r + 1];
		new_seed *rec = w.rec + b;
		int32_t *recp = w.recp + b;
		int n_new = 0;
		int t_next = b + lane < e ? w.t[b + lane] : -1;                          // mark word of the tile ahead (decides which loads follow)
		for (int base = b; base < e; base += 32) {
			const int i = base + lane;
			const bool valid = i < e;
			int pi = -1, vi = 0, fi = 0;
			const int ti = t_next;
			t_next = i + 32 < e ? w.t[i + 32] : -1;
			if (valid) {
				if (ti <= -2) fi = vi = -2 - ti;                                  // singleton: k_prep left only its span
				else { const int4 ri = w.wrec[i]; pi = ri.w; fi = ri.z; vi = w.v[i]; }
			}
			const bool E = valid && (vi >= min_sc || pi >= 0);                  // chain.c:109
			// chain.c:93-107: a predecessor that was not emitted on its own is emitted by its FIRST referencer.  Anchors are
			// visited in increasing i, so "first" = no earlier tile emitted it (newid still -1; a predecessor inside this tile
			// cannot have been emitted earlier, its referencers all sit in this tile or later) and the lowest lane of this
			// tile among those that name it.
			bool cand = false;
			int vp = 0, fp = 0, nid_p = -1;
			if (pi >= 0) {
				const int4 rp = w.wrec[pi];
				vp = w.v[pi]; fp = rp.z;
				if (pi < base) nid_p = w.newid[pi];
				cand = !(vp >= min_sc || rp.w >= 0) && nid_p < 0;
			}
			// (most tiles have no candidate at all -- only a chain's first anchor is one -- and skip the vote; lanes without a
			// candidate share one group, MATCH.ANY gets slower with the number of distinct values)
			unsigned grp = 0;
			if (__any_sync(FULL, cand)) grp = __match_any_sync(FULL, cand ? pi : -1);
			const int first = __ffs(grp) - 1;                                   // first referencer of this lane's candidate
			const bool late = cand && first == lane;
			// records emitted by the lanes below this one: two ballots instead of a shuffle scan (a lane emits 0, 1 or 2)
			const unsigned em = __ballot_sync(FULL, E), lm = __ballot_sync(FULL, late);
			const unsigned lt = lanemask_lt(lane);
			const int slot = n_new + __popc(em & lt) + __popc(lm & lt);          // first record this lane emits
			const int my_new = slot + (late ? 1 : 0);
			// new index of the predecessor without a trip through memory: emitted by this tile's first referencer (its slot),
			// emitted on its own inside this tile (that lane's my_new), or known from an earlier tile (nid_p)
			const int pn_late = __shfl_sync(FULL, slot, first & 31);
			const int pn_tile = __shfl_sync(FULL, my_new, (pi - base) & 31);
			const int pn = pi < 0 ? -1 : (cand ? pn_late : (pi >= base ? pn_tile : nid_p));
			if (valid) w.newid[i] = E ? my_new : -1;
			__syncwarp();
			if (late) {
				w.newid[pi] = slot;
				const int32_t pw = (int32_t)(0xfffffffcu | (vp >= min_sc ? 1u : 0u) | (fp < vp ? 2u : 0u));
				store_rec(&rec[slot], w.a[pi], pw, fp);
				recp[slot] = pw;
			}
			if (E) {
				const int32_t pw = (int32_t)(((unsigned)pn << 2) | (vi >= min_sc ? 1u : 0u) | (fi < vi ? 2u : 0u));
				store_rec(&rec[my_new], w.a[i], pw, fi);
				recp[my_new] = pw;
			}
			__syncwarp();                                                       // newid[] of this tile is visible to the next one
			n_new += __popc(em) + __popc(lm);
		}
		if (lane == 0) w.new_i[r] = (uint32_t)n_new;
	}
}

// bottom_batch: records come from the host; derive recp[] and new_i[]
__global__ void __launch_bounds__(256) k_split_records(Workspace w)
{
	const int n = w.n_anchors;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) w.recp[i] = w.rec[i].p;
	for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < w.n_reads; r += gridDim.x * blockDim.x)
		w.new_i[r] = (uint32_t)(w.offs[r + 1] - w.offs[r]);
}

// ---------------------------------------------------------------------------------------------------------
// k_tail: chain.c:127-214 for one read per CTA.
// ---------------------------------------------------------------------------------------------------------
#define TAIL_T 128

// ascending in-place sort of z[0..n) by the whole CTA ("flip" bitonic network: every compare-exchange is
// ascending, so a virtual +inf padding beyond n never has to move)
static __device__ void cta_sort_u64(uint64_t *z, int n)
{
	const int tid = threadIdx.x;
	for (int k = 2; (k >> 1) < n; k <<= 1) {
		for (int i = tid; i < n; i += TAIL_T) {
			const int l = i ^ (k - 1);
			if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
		}
		__syncthreads();
		for (int j = k >> 2; j > 0; j >>= 1) {
			for (int i = tid; i < n; i += TAIL_T) {
				const int l = i ^ j;
				if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
			}
			__syncthreads();
		}
	}
}

// ascending in-place sort of z[0..n) by one warp (same flip-bitonic network as cta_sort_u64)
static __device__ void warp_sort_u64(uint64_t *z, int n, int lane)
{
	for (int k = 2; (k >> 1) < n; k <<= 1) {
		for (int i = lane; i < n; i += 32) {
			const int l = i ^ (k - 1);
			if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
		}
		__syncwarp();
		for (int j = k >> 2; j > 0; j >>= 1) {
			for (int i = lane; i < n; i += 32) {
				const int l = i ^ j;
				if (l > i && l < n) { const uint64_t x = z[i], y = z[l]; if (x > y) { z[i] = y; z[l] = x; } }
			}
			__syncwarp();
		}
	}
}

// ---------------------------------------------------------------------------------------------------------
// k_tail_warp: chain.c:127-214 for reads with at most TW_CAP records, one WARP per read.  The predecessor words
// and the successor / claim flags of the read are staged in shared memory, so the two pointer walks of the
// reference (end -> peak, chain.c:151; peak -> chain start, chain.c:167-172) cost one LDS per step instead of a
// dependent global load.  Larger reads are left to k_tail (CTA per read, global memory).
// ---------------------------------------------------------------------------------------------------------
#define TW_CAP 1024
#define TW_WARPS 8

__global__ void __launch_bounds__(TW_WARPS * 32) k_tail_warp(Workspace w, int32_t min_cnt, int32_t min_sc)
{
	__shared__ int s_rp_all[TW_WARPS][TW_CAP];          // rec[].p words; reused as head/tail of the radix replay
	__shared__ uint8_t s_flag_all[TW_WARPS][TW_CAP];    // "has a successor", then "claimed"
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	int *s_rp = s_rp_all[wib];
	uint8_t *s_flag = s_flag_all[wib];
	const int gw = blockIdx.x * TW_WARPS + wib, nw = gridDim.x * TW_WARPS;
	for (int r = gw; r < w.n_reads; r += nw) {
		const int m = (int)w.new_i[r];
		if (m > TW_CAP) continue;                         // k_tail's job
		if (m == 0) { if (lane == 0) { w.n_u[r] = 0; w.n_v[r] = 0; } continue; }
		const int base = (int)w.offs[r];
		const int cap = (int)(w.offs[r + 1] - w.offs[r]);
		const new_seed *rec = w.rec + base;
		int32_t *members = w.members + base, *voff = w.voff + base, *order = w.order + base, *ooff = w.ooff + base;
		uint64_t *cand = w.cand + base, *chain_u = w.chain_u + base, *wy = w.wy + base;

		__syncwarp();
		for (int i = lane; i < m; i += 32) { s_rp[i] = w.recp[base + i]; s_flag[i] = 0; }
		__syncwarp();
		// chain.c:128-131: who is somebody's predecessor
		for (int i = lane; i < m; i += 32) { const int pp = s_rp[i]; if (pp >= 0) s_flag[pp >> 2] = 1; }
		__syncwarp();
		// chain.c:132-155: ends (bit0 set, no successor) -> peak (walk while bit1 set)
		int n_end = 0;
		for (int i0 = 0; i0 < m; i0 += 32) {
			const int i = i0 + lane;
			bool is_end = false;
			int pp = 0;
			if (i < m) { pp = s_rp[i]; is_end = (pp & 1) && !s_flag[i]; }
			const unsigned em = __ballot_sync(FULL, is_end);
			if (is_end) {
				int j = i;
				while (j >= 0 && (pp & 2)) { j = pp >> 2; if (j >= 0) pp = s_rp[j]; }
				if (j < 0) j = i;
				cand[n_end + __popc(em & lanemask_lt(lane))] = ((uint64_t)(uint32_t)rec[j].f << 32) | (uint32_t)j;
			}
			n_end += __popc(em);
		}
		if (n_end == 0) { if (lane == 0) { w.n_u[r] = 0; w.n_v[r] = 0; } continue; }
		__syncwarp();
		// chain.c:156-160: descending (peak score, peak index) == ascending sort read backwards
		if (n_end > 1) warp_sort_u64(cand, n_end, lane);
		for (int i = lane; i < m; i += 32) s_flag[i] = 0;        // from here on: claim marks
		__syncwarp();

		// chain.c:163-179: backtrack in that order; claims persist even for rejected chains.  The walk is sequential by
		// definition, but most chains link consecutive records (p[k] == k-1), so the warp takes a whole run of
		// consecutively linked, unclaimed records per iteration: lane l looks at record j-l, the run ends at the first
		// lane whose record is not the predecessor of the previous one (or is already claimed).
		int n_u = 0, n_v = 0;
		for (int ii = n_end - 1; ii >= 0; --ii) {
			const uint64_t c = cand[ii];
			const int peak_sc = (int)(c >> 32);
			int j = (int)(uint32_t)c;
			const int n0 = n_v;
			for (;;) {
				const int cl = j - lane;
				const int wl = cl >= 0 ? s_rp[cl] : -4;
				const bool claimed_l = cl >= 0 ? (s_flag[cl] != 0) : true;
				const int wprev = __shfl_up_sync(FULL, wl, 1);
				// lane 0 is always taken: j is unclaimed by construction, except the peak, which the reference takes anyway (do-while)
				const bool ok = lane == 0 ? true : ((wprev >> 2) == cl && cl >= 0 && !claimed_l);
				const int run = __ffs(~__ballot_sync(FULL, ok)) - 1;   // >= 1; -1 means all 32 lanes link
				const int take = run < 0 ? 32 : run;
				if (lane < take) {
					if (n_v + lane < cap) members[n_v + lane] = cl;
					s_flag[cl] = 1;
				}
				n_v += take;
				const int jn = __shfl_sync(FULL, wl, take - 1) >> 2;   // predecessor of the last record taken
				__syncwarp();
				j = jn;
				if (jn < 0 || s_flag[jn]) break;
			}
			const int cnt = n_v - n0;
			bool keep = false;
			uint64_t val = 0;
			if (n_v > cap) { /* would overrun the reference's v[] as well (chain.c:168); drop */ }
			else if (j < 0) {
				if (cnt >= min_cnt) { keep = true; val = ((uint64_t)(uint32_t)peak_sc << 32) | (uint32_t)cnt; }
			} else {
				const int d = peak_sc - rec[j].f;
				if (d >= min_sc && cnt >= min_cnt) { keep = true; val = ((uint64_t)(uint32_t)d << 32) | (uint32_t)cnt; }
			}
			if (keep) { if (lane == 0) { chain_u[n_u] = val; voff[n_u] = n0; } ++n_u; }
			else n_v = n0;
		}
		if (lane == 0) { w.n_u[r] = n_u; w.n_v[r] = n_v; }
		if (n_u == 0) continue;
		__syncwarp();
		// chain.c:199-204: w[k] = { x of the chain's first anchor, offset<<32 | k }, sorted by x only
		uint64_t *wx = cand;
		for (int k = lane; k < n_u; k += 32) {
			const int cnt = (int)(uint32_t)chain_u[k];
			wx[k] = rec[members[voff[k] + cnt - 1]].seed.x;
			wy[k] = ((uint64_t)(uint32_t)voff[k] << 32) | (uint32_t)k;
		}
		__syncwarp();
		if (n_u <= RS_MIN) {
			// insertion sort is stable: rank = #{smaller} + #{equal and earlier}
			for (int k = lane; k < n_u; k += 32) {
				const uint64_t xk = wx[k];
				int rank = 0;
				for (int k2 = 0; k2 < n_u; ++k2) { const uint64_t x2 = wx[k2]; rank += (x2 < xk) || (x2 == xk && k2 < k); }
				order[rank] = k;
			}
		} else {
			if (lane == 0) radix_sort_kv_exact(wx, wy, n_u, w.stk + base, s_rp, s_rp + 256);
			__syncwarp();
			for (int k = lane; k < n_u; k += 32) order[k] = (int)(uint32_t)wy[k];
		}
		__syncwarp();
		if (lane == 0) {
			int acc = 0;
			for (int i = 0; i < n_u; ++i) { ooff[i] = acc; acc += (int)(uint32_t)chain_u[order[i]]; }
		}
	}
}

__global__ void __launch_bounds__(TAIL_T) k_tail(Workspace w, int32_t min_cnt, int32_t min_sc, int32_t min_m)
{
	__shared__ int s_cnt, s_nu;
	__shared__ int s_head[256], s_tail[256];
	const int tid = threadIdx.x;
	for (int r = blockIdx.x; r < w.n_reads; r += gridDim.x) {
		const int base = (int)w.offs[r];
		const int cap = (int)(w.offs[r + 1] - w.offs[r]);
		const int m = (int)w.new_i[r];
		if (m <= min_m) continue;                           // k_tail_warp's job (it also handles m == 0)
		const int32_t *rp = w.recp + base;
		const new_seed *rec = w.rec + base;
		uint8_t *succ = w.succ + base, *claimed = w.claimed + base;
		int32_t *members = w.members + base, *voff = w.voff + base, *order = w.order + base, *ooff = w.ooff + base;
		uint64_t *cand = w.cand + base, *chain_u = w.chain_u + base, *wy = w.wy + base;

		// chain.c:128-131: who is somebody's predecessor
		for (int i = tid; i < m; i += TAIL_T) { succ[i] = 0; claimed[i] = 0; }
		if (tid == 0) s_cnt = 0;
		__syncthreads();
		for (int i = tid; i < m; i += TAIL_T) { const int pp = rp[i]; if (pp >= 0) succ[pp >> 2] = 1; }
		__syncthreads();
		// chain.c:132-155: ends (bit0 set, no successor) -> peak (walk while bit1 set)
		for (int i = tid; i < m; i += TAIL_T) {
			int pp = rp[i];
			if ((pp & 1) && !succ[i]) {
				int j = i;
				while (j >= 0 && (pp & 2)) { j = pp >> 2; if (j >= 0) pp = rp[j]; }
				if (j < 0) j = i;
				cand[atomicAdd(&s_cnt, 1)] = ((uint64_t)(uint32_t)rec[j].f << 32) | (uint32_t)j;
			}
		}
		__syncthreads();
		const int n_end = s_cnt;
		if (n_end == 0) {
			if (tid == 0) { w.n_u[r] = 0; w.n_v[r] = 0; }
			__syncthreads();
			continue;
		}
		// chain.c:156-160: descending (peak score, peak index) == ascending sort read backwards
		cta_sort_u64(cand, n_end);

		// chain.c:163-179: backtrack in that order; claims persist even for rejected chains.  Warp 0 walks run-wise like
		// k_tail_warp (a stretch of consecutively linked, unclaimed records per iteration), here on global memory.
		if (tid < 32) {
			const int lane = tid;
			int n_v = 0, k = 0;
			for (int ii = n_end - 1; ii >= 0; --ii) {
				const uint64_t c = cand[ii];
				const int peak_sc = (int)(c >> 32);
				int j = (int)(uint32_t)c;
				const int n0 = n_v;
				for (;;) {
					const int cl = j - lane;
					const int wl = cl >= 0 ? rp[cl] : -4;
					const bool claimed_l = cl >= 0 ? (claimed[cl] != 0) : true;
					const int wprev = __shfl_up_sync(FULL, wl, 1);
					const bool ok = lane == 0 ? true : ((wprev >> 2) == cl && cl >= 0 && !claimed_l);
					const int run = __ffs(~__ballot_sync(FULL, ok)) - 1;
					const int take = run < 0 ? 32 : run;
					if (lane < take) {
						if (n_v + lane < cap) members[n_v + lane] = cl;
						claimed[cl] = 1;
					}
					n_v += take;
					const int jn = __shfl_sync(FULL, wl, take - 1) >> 2;
					__syncwarp();
					j = jn;
					if (jn < 0 || claimed[jn]) break;
				}
				const int cnt = n_v - n0;
				bool keep = false;
				uint64_t val = 0;
				if (n_v > cap) { /* would overrun the reference's v[] as well (chain.c:168); drop */ }
				else if (j < 0) {
					if (cnt >= min_cnt) { keep = true; val = ((uint64_t)(uint32_t)peak_sc << 32) | (uint32_t)cnt; }
				} else {
					const int d = peak_sc - rec[j].f;
					if (d >= min_sc && cnt >= min_cnt) { keep = true; val = ((uint64_t)(uint32_t)d << 32) | (uint32_t)cnt; }
				}
				if (keep) { if (lane == 0) { chain_u[k] = val; voff[k] = n0; } ++k; }
				else n_v = n0;
			}
			if (lane == 0) { s_nu = k; w.n_u[r] = k; w.n_v[r] = n_v; }
		}
		__syncthreads();
		const int n_u = s_nu;
		if (n_u > 0) {
			// chain.c:199-204: w[k] = { x of the chain's first anchor, offset<<32 | k }, sorted by x only
			uint64_t *wx = cand;   // candidates are consumed; reuse as sort keys
			for (int k = tid; k < n_u; k += TAIL_T) {
				const int cnt = (int)(uint32_t)chain_u[k];
				wx[k] = rec[members[voff[k] + cnt - 1]].seed.x;
				wy[k] = ((uint64_t)(uint32_t)voff[k] << 32) | (uint32_t)k;
			}
			__syncthreads();
			if (n_u <= RS_MIN) {
				// insertion sort is stable: rank = #{smaller} + #{equal and earlier}
				for (int k = tid; k < n_u; k += TAIL_T) {
					const uint64_t xk = wx[k];
					int rank = 0;
					for (int k2 = 0; k2 < n_u; ++k2) { const uint64_t x2 = wx[k2]; rank += (x2 < xk) || (x2 == xk && k2 < k); }
					order[rank] = k;
				}
			} else {
				if (tid == 0) radix_sort_kv_exact(wx, wy, n_u, w.stk + base, s_head, s_tail);
				__syncthreads();
				for (int k = tid; k < n_u; k += TAIL_T) order[k] = (int)(uint32_t)wy[k];
			}
			__syncthreads();
			if (tid == 0) {
				int acc = 0;
				for (int i = 0; i < n_u; ++i) { ooff[i] = acc; acc += (int)(uint32_t)chain_u[order[i]]; }
			}
		}
		__syncthreads();
	}
}

// exclusive scans over reads: n_u -> u_off, n_v -> b_off, new_i -> rec_off; totals.  Three small launches (per-block
// sums, scan of the block sums, per-block scan + offset) so that every access is coalesced and all SMs take part.
static __device__ __forceinline__ void block_scan3(long long (&x)[3], long long (&total)[3], long long (*s_part)[32])
{	// inclusive scan of x[] across the 1024 threads of the block; total[] = block sums
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
	for (int q = 0; q < 3; ++q) {
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { const long long o = __shfl_up_sync(FULL, x[q], d); if (lane >= d) x[q] += o; }
		if (lane == 31) s_part[q][wid] = x[q];
	}
	__syncthreads();
	if (wid == 0) {
#pragma unroll
		for (int q = 0; q < 3; ++q) {
			long long y = s_part[q][lane];
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { const long long o = __shfl_up_sync(FULL, y, d); if (lane >= d) y += o; }
			s_part[q][lane] = y;
		}
	}
	__syncthreads();
#pragma unroll
	for (int q = 0; q < 3; ++q) { if (wid) x[q] += s_part[q][wid - 1]; total[q] = s_part[q][31]; }
}

static __device__ __forceinline__ void load_read_counts(const Workspace &w, int r, int have_tail, long long (&val)[3])
{
	val[0] = val[1] = val[2] = 0;
	if (r < w.n_reads) {
		if (have_tail) { val[0] = w.n_u[r]; val[1] = w.n_v[r]; }
		val[2] = w.new_i[r];
	}
}

__global__ void __launch_bounds__(1024) k_scan_block_sums(Workspace w, int have_tail)
{
	__shared__ long long s_part[3][32];
	long long val[3], tot[3];
	load_read_counts(w, blockIdx.x * 1024 + threadIdx.x, have_tail, val);
	block_scan3(val, tot, s_part);
	if (threadIdx.x == 0) { w.scan_part[3 * blockIdx.x] = tot[0]; w.scan_part[3 * blockIdx.x + 1] = tot[1]; w.scan_part[3 * blockIdx.x + 2] = tot[2]; }
}

__global__ void __launch_bounds__(1024) k_scan_block_offsets(Workspace w, int n_blocks)
{	// one CTA: exclusive scan of the block sums in place; totals
	__shared__ long long s_part[3][32];
	__shared__ long long s_carry[3];
	if (threadIdx.x < 3) s_carry[threadIdx.x] = 0;
	__syncthreads();
	for (int base = 0; base < n_blocks; base += 1024) {
		const int c = base + threadIdx.x;
		long long val[3] = {0, 0, 0}, x[3], tot[3];
		if (c < n_blocks) { val[0] = w.scan_part[3 * c]; val[1] = w.scan_part[3 * c + 1]; val[2] = w.scan_part[3 * c + 2]; }
		x[0] = val[0]; x[1] = val[1]; x[2] = val[2];
		block_scan3(x, tot, s_part);
		if (c < n_blocks) {
#pragma unroll
			for (int q = 0; q < 3; ++q) w.scan_part[3 * c + q] = s_carry[q] + x[q] - val[q];
		}
		__syncthreads();
		if (threadIdx.x < 3) s_carry[threadIdx.x] += tot[threadIdx.x];
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		w.u_off[w.n_reads] = s_carry[0]; w.b_off[w.n_reads] = s_carry[1]; w.rec_off[w.n_reads] = s_carry[2];
		w.totals[0] = s_carry[0]; w.totals[1] = s_carry[1]; w.totals[2] = w.counters[CNT_UNITS] + w.counters[CNT_LONG_UNITS] + w.counters[CNT_N + CNT_UNITS] + w.counters[CNT_N + CNT_LONG_UNITS];   // (+ second half, see chain_device)
		w.totals[3] = s_carry[2];
		w.totals[4] = (long long)w.visited[0]; w.totals[5] = (long long)w.visited[1];
	}
}

__global__ void __launch_bounds__(1024) k_scan_apply(Workspace w, int have_tail)
{
	__shared__ long long s_part[3][32];
	const int r = blockIdx.x * 1024 + threadIdx.x;
	long long val[3], x[3], tot[3];
	load_read_counts(w, r, have_tail, val);
	x[0] = val[0]; x[1] = val[1]; x[2] = val[2];
	block_scan3(x, tot, s_part);
	if (r < w.n_reads) {
		w.u_off[r] = w.scan_part[3 * blockIdx.x] + x[0] - val[0];
		w.b_off[r] = w.scan_part[3 * blockIdx.x + 1] + x[1] - val[1];
		w.rec_off[r] = w.scan_part[3 * blockIdx.x + 2] + x[2] - val[2];
	}
}

static int launch_scan_reads(const Workspace &w, int have_tail, cudaStream_t s)
{
	const int nb = (w.n_reads + 1023) / 1024 > 0 ? (w.n_reads + 1023) / 1024 : 1;
	k_scan_block_sums<<<nb, 1024, 0, s>>>(w, have_tail);
	k_scan_block_offsets<<<1, 1024, 0, s>>>(w, nb);
	k_scan_apply<<<nb, 1024, 0, s>>>(w, have_tail);
	return 3;
}

// chain.c:187-192,205-213: write the final packed u[] and b[]
__global__ void __launch_bounds__(TAIL_T) k_gather(Workspace w)
{
	const int tid = threadIdx.x;
	for (int r = blockIdx.x; r < w.n_reads; r += gridDim.x) {
		const int n_u = w.n_u[r];
		if (n_u == 0) continue;
		const int base = (int)w.offs[r];
		const long long ub = w.u_off[r], bb = w.b_off[r];
		const new_seed *rec = w.rec + base;
		const int32_t *members = w.members + base;
		for (int i = 0; i < n_u; ++i) {
			const int src = w.order[base + i];
			const uint64_t cu = w.chain_u[base + src];
			const int cnt = (int)(uint32_t)cu, vo = w.voff[base + src], oo = w.ooff[base + i];
			if (tid == 0) w.u_out[ub + i] = cu;
			for (int k = tid; k < cnt; k += TAIL_T) {
				const new_seed *s = &rec[members[vo + cnt - 1 - k]];
				w.b_out[bb + oo + k] = make_ulonglong2(s->seed.x, s->seed.y);
			}
		}
	}
}

// fill_batch: pack the per-read record arrays (chain.c:324-326 surface)
__global__ void __launch_bounds__(TAIL_T) k_pack_records(Workspace w)
{
	const int tid = threadIdx.x;
	for (int r = blockIdx.x; r < w.n_reads; r += gridDim.x) {
		const int m = (int)w.new_i[r];
		const unsigned long long *src = reinterpret_cast<const unsigned long long*>(w.rec + w.offs[r]);
		unsigned long long *dst = reinterpret_cast<unsigned long long*>(w.rec_out + w.rec_off[r]);
		for (int k = tid; k < 3 * m; k += TAIL_T) dst[k] = src[k];
	}
}

// debug: export f[] and p[] (per-read indices, what the reference's arrays hold at chain.c:89) from the records
__global__ void __launch_bounds__(256) k_export_fp(Workspace w)
{
	const int lane = threadIdx.x & 31;
	const int gw = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
	const int nw = (int)((gridDim.x * (unsigned)blockDim.x) >> 5);
	for (int r = gw; r < w.n_reads; r += nw) {
		const int b = (int)w.offs[r], e = (int)w.offs[r + 1];
		for (int i = b + lane; i < e; i += 32) {
			const int ti = w.t[i];
			if (ti <= -2) { w.f[i] = -2 - ti; w.p[i] = -1; w.v[i] = -2 - ti; continue; }   // singleton (see k_prep)
			const int4 ri = w.wrec[i];
			w.f[i] = ri.z;
			w.p[i] = ri.w >= 0 ? ri.w - b : -1;
		}
	}
}

// ---------------------------------------------------------------------------------------------------------
// Integer-issue microbenchmark: the denominator of the predecessor-scan roofline.  Eight
// independent dependency chains per thread; MODE 0 = ALU pipe only (add/xor), MODE 1 = ALU + FMA-pipe integer
// (mad.lo) interleaved, which is what can saturate the 1 instruction/clk/SMSP issue port.
// ---------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) k_int_peak(unsigned *out, int iters, unsigned seed)
{
	unsigned a0 = threadIdx.x + seed, a1 = a0 * 3u + 1u, a2 = a0 * 5u + 2u, a3 = a0 * 7u + 3u;
	unsigned a4 = a0 * 11u + 4u, a5 = a0 * 13u + 5u, a6 = a0 * 17u + 6u, a7 = a0 * 19u + 7u;
	const unsigned c = seed | 1u, d = seed * 2654435761u + 12345u;
	for (int it = 0; it < iters; ++it) {
#pragma unroll
		for (int k = 0; k < 8; ++k) {
			if (MODE == 0) {
				// every chain alternates add / xor so that ptxas cannot fold two steps into one IADD3 / LOP3
				if (k & 1) {
					asm volatile("add.u32 %0, %0, %1;" : "+r"(a0) : "r"(c));  asm volatile("xor.b32 %0, %0, %1;" : "+r"(a1) : "r"(d));
					asm volatile("add.u32 %0, %0, %1;" : "+r"(a2) : "r"(c));  asm volatile("xor.b32 %0, %0, %1;" : "+r"(a3) : "r"(d));
					asm volatile("add.u32 %0, %0, %1;" : "+r"(a4) : "r"(c));  asm volatile("xor.b32 %0, %0, %1;" : "+r"(a5) : "r"(d));
					asm volatile("add.u32 %0, %0, %1;" : "+r"(a6) : "r"(c));  asm volatile("xor.b32 %0, %0, %1;" : "+r"(a7) : "r"(d));
				} else {
					asm volatile("xor.b32 %0, %0, %1;" : "+r"(a0) : "r"(d));  asm volatile("add.u32 %0, %0, %1;" : "+r"(a1) : "r"(c));
					asm volatile("xor.b32 %0, %0, %1;" : "+r"(a2) : "r"(d));  asm volatile("add.u32 %0, %0, %1;" : "+r"(a3) : "r"(c));
					asm volatile("xor.b32 %0, %0, %1;" : "+r"(a4) : "r"(d));  asm volatile("add.u32 %0, %0, %1;" : "+r"(a5) : "r"(c));
					asm volatile("xor.b32 %0, %0, %1;" : "+r"(a6) : "r"(d));  asm volatile("add.u32 %0, %0, %1;" : "+r"(a7) : "r"(c));
				}
			} else {
				asm volatile("add.u32 %0, %0, %1;" : "+r"(a0) : "r"(c));  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a1) : "r"(c), "r"(d));
				asm volatile("xor.b32 %0, %0, %1;" : "+r"(a2) : "r"(d));  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a3) : "r"(c), "r"(d));
				asm volatile("add.u32 %0, %0, %1;" : "+r"(a4) : "r"(c));  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a5) : "r"(c), "r"(d));
				asm volatile("xor.b32 %0, %0, %1;" : "+r"(a6) : "r"(d));  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a7) : "r"(c), "r"(d));
			}
		}
	}
	const unsigned r = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
	if (r == 0x12345678u) out[0] = r;   // keep the chains alive
}

// returns lane-ops issued by one launch (grid * 256 threads * iters * 64)
long long launch_int_peak(int mode, int grid, int iters, unsigned *d_out, cudaStream_t s)
{
	if (mode == 0) k_int_peak<0><<<grid, 256, 0, s>>>(d_out, iters, 77u);
	else k_int_peak<1><<<grid, 256, 0, s>>>(d_out, iters, 77u);
	return (long long)grid * 256 * (long long)iters * 64;
}

// ---------------------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------------------
static inline int blocks_for(long long items, int per_block, int cap)
{
	long long b = (items + per_block - 1) / per_block;
	if (b < 1) b = 1;
	if (b > cap) b = cap;
	return (int)b;
}

template <int RING> static int fill_smem_bytes(int tab_entries) { return FillSmem<RING>::tab + FILL_WARPS * tab_entries * 2 + RING * 16; }   // + alignment slack
static int fill_tab_entries(const mm2b200_params_t &prm)
{	// 16-bit table of bw+1 gap costs per warp, only while it stays small (<= 2 KB per warp) and the costs fit
	if (prm.bw < 0 || prm.bw > 1023) return 0;
	return (prm.bw + 1 + 7) & ~7;
}

static int g_fill_ctas_cap = 0;   // MM2B200_FILL_CTAS: cap on resident fill CTAs per SM (0 = occupancy), see tools/overlap_probe.py
// per-device occupancy cache of one kernel instance (a host thread may drive several devices)
struct OccCache { int dev = -1, smem = -1, occ = 0; };
template <typename K>
static int cached_occupancy(OccCache &oc, K kern, int threads, int smem)
{
	int dev = 0;
	cudaGetDevice(&dev);
	if (oc.dev != dev || oc.smem != smem) {
		cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
		int occ = 0;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem);
		oc.occ = occ < 1 ? 1 : occ; oc.smem = smem; oc.dev = dev;
	}
	return oc.occ;
}

template <bool TABLE, bool FULLV, bool WIDE, int RING, bool DEEP>
static void launch_fill_ring(const Workspace &w, const mm2b200_params_t &prm, int sms, int tab_entries, cudaStream_t s)
{
	const int smem = fill_smem_bytes<RING>(tab_entries);
	auto kern = k_fill_fast<TABLE, FULLV, WIDE, RING, DEEP>;
	static thread_local OccCache oc;   // one cache per template instance
	const int occ = cached_occupancy(oc, kern, FILL_WARPS * 32, smem);
	const int per_sm = (g_fill_ctas_cap > 0 && g_fill_ctas_cap < occ) ? g_fill_ctas_cap : occ;
	kern<<<sms * per_sm, FILL_WARPS * 32, smem, s>>>(w, prm, tab_entries);
}
template <bool TABLE, bool FULLV, bool WIDE>
static void launch_fill_fast(const Workspace &w, const mm2b200_params_t &prm, int sms, int tab_entries, int ring, cudaStream_t s)
{
	// ring < 0: the long-read variant of the 64-anchor ring (DEEP: the global part of a window four chunks at a time; 64
	// registers, 8 CTAs per SM).  Deeper rings exist for experiments (MM2B200_RING) and always use the batched loads.
	if (ring >= 256) launch_fill_ring<TABLE, FULLV, WIDE, 256, true>(w, prm, sms, tab_entries, s);
	else if (ring >= 128) launch_fill_ring<TABLE, FULLV, WIDE, 128, true>(w, prm, sms, tab_entries, s);
	else if (ring < 0) launch_fill_ring<TABLE, FULLV, WIDE, 64, true>(w, prm, sms, tab_entries, s);
	else launch_fill_ring<TABLE, FULLV, WIDE, 64, false>(w, prm, sms, tab_entries, s);
}

void query_launch_cfg(LaunchCfg *cfg)
{
	int dev = 0, sms = 148, occ = 0;
	cudaGetDevice(&dev);
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_fill_generic, 128, 0);
	if (occ < 1) occ = 1;
	cfg->sm_count = sms;
	cfg->fill_blocks_per_sm = occ;
	const char *ec = getenv("MM2B200_FILL_CTAS");
	g_fill_ctas_cap = ec ? atoi(ec) : 0;
	const char *er = getenv("MM2B200_RING");       // force the fill kernel's ring depth (64, 128, 256); default: by read length
	cfg->fill_ring = er ? atoi(er) : 0;
}

int launch_prep(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s)
{
	// a whole batch clears both counter sets; a half of a split device batch only its own (the halves run on two streams,
	// and the caller clears visited[] before it forks)
	cudaMemsetAsync(w.counters, 0, (w.half ? 1 : 2) * CNT_N * sizeof(int32_t), s);
	if (!w.half) cudaMemsetAsync(w.visited, 0, 2 * sizeof(unsigned long long), s);
	k_prep<<<blocks_for(w.n_reads, 8, 148 * 64), 256, 0, s>>>(w, prm.max_dist_x);
	return 1;
}

int launch_fill(const Workspace &w, const mm2b200_params_t &prm, const LaunchCfg &cfg, bool host_generic, cudaStream_t s)
{
	int n = 0;
	// Both variants are enqueued unless the host already knows the generic one is needed; the one that does not
	// match the device-side segment flag returns immediately.
	if (!host_generic) {
		const int te = fill_tab_entries(prm);
		const bool full = cfg.count_visited || prm.max_iter > 0;   // instrumented / max_iter-aware variant
		// Long reads (thousands of anchors per read: map-pb, asm5, ava) scan hundreds of predecessors per anchor, far beyond the
		// first chunks, and every chunk out of the ring is an L2 / HBM round trip: those sub-batches run the variants with the
		// batched deep loads.  Which ring depth depends on how full the GPU gets (measured, tools/ring_sweep.sh): with fewer
		// runs than resident warps (<= 2048 reads) the 256-anchor ring wins (fewer trips, occupancy is not the limit); with
		// more, the 64-anchor ring (8 CTAs per SM instead of 5).  Ordinary reads rarely leave the first two chunks and keep
		// the lean 64-anchor variant (48 registers, 10 CTAs per SM).
		const bool long_reads = w.n_reads > 0 && w.n_anchors / w.n_reads >= 4096;
		const int ring = cfg.fill_ring ? cfg.fill_ring : (long_reads ? (w.n_reads <= 2048 ? 256 : -1) : 64);
		if (prm.max_dist_x >= (1 << 26)) {                         // per-lane window test not exact (33 gaps could wrap 2^32): general chunk routine only
			if (te) launch_fill_fast<true, true, true>(w, prm, cfg.sm_count, te, ring, s); else launch_fill_fast<false, true, true>(w, prm, cfg.sm_count, 0, ring, s);
		} else if (te) { if (full) launch_fill_fast<true, true, false>(w, prm, cfg.sm_count, te, ring, s); else launch_fill_fast<true, false, false>(w, prm, cfg.sm_count, te, ring, s); }
		else { if (full) launch_fill_fast<false, true, false>(w, prm, cfg.sm_count, 0, ring, s); else launch_fill_fast<false, false, false>(w, prm, cfg.sm_count, 0, ring, s); }
		++n;
	}
	k_fill_generic<<<cfg.sm_count * cfg.fill_blocks_per_sm, 128, 0, s>>>(w, prm); ++n;
	return n;
}

int launch_compact(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s)
{
	k_compact<<<blocks_for(w.n_reads, 8, 148 * 64), 256, 0, s>>>(w, prm.min_sc);
	return 1;
}

int launch_split_records(const Workspace &w, cudaStream_t s)
{
	k_split_records<<<blocks_for(w.n_anchors > w.n_reads ? w.n_anchors : w.n_reads, 256 * 4, 148 * 16), 256, 0, s>>>(w);
	return 1;
}

// the two halves of launch_tail, for callers that run the per-read part of several read ranges on different streams and
// the scan over all reads once
int launch_tail_reads(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s)
{
	k_tail_warp<<<blocks_for(w.n_reads, TW_WARPS, 148 * 16), TW_WARPS * 32, 0, s>>>(w, prm.min_cnt, prm.min_sc);
	k_tail<<<blocks_for(w.n_reads, 1, 148 * 64), TAIL_T, 0, s>>>(w, prm.min_cnt, prm.min_sc, TW_CAP);
	return 2;
}
int launch_tail_scan(const Workspace &w, cudaStream_t s) { return launch_scan_reads(w, 1, s); }

int launch_tail(const Workspace &w, const mm2b200_params_t &prm, cudaStream_t s)
{
	k_tail_warp<<<blocks_for(w.n_reads, TW_WARPS, 148 * 16), TW_WARPS * 32, 0, s>>>(w, prm.min_cnt, prm.min_sc);
	k_tail<<<blocks_for(w.n_reads, 1, 148 * 64), TAIL_T, 0, s>>>(w, prm.min_cnt, prm.min_sc, TW_CAP);
	return 2 + launch_scan_reads(w, 1, s);
}

int launch_gather(const Workspace &w, cudaStream_t s)
{
	k_gather<<<blocks_for(w.n_reads, 1, 148 * 64), TAIL_T, 0, s>>>(w);
	return 1;
}

int launch_pack_records(const Workspace &w, cudaStream_t s)
{
	const int n = launch_scan_reads(w, 0, s);
	k_pack_records<<<blocks_for(w.n_reads, 1, 148 * 64), TAIL_T, 0, s>>>(w);
	return n + 1;
}

// ---------------------------------------------------------------------------------------------------------
// wire format of the upload (host batcher -> device): 8 bytes per anchor plus small dictionaries instead of 16.
// Inside a block of WIRE_BLOCK consecutive anchors the high half of x (strand|rid) and the high half of y
// (flags|seg|qspan, minimap.h:48-55) take few distinct values, so each block carries one dictionary for each and an
// anchor travels as  low32(x) | (qpos | iy << qbits) << 32 | ix << 56  with qbits = 24 - ybits, ybits = bits needed for
// the block's y dictionary.  Block descriptor: dictionary position (32) | x entries (16) | ybits (16); the y entries
// follow the x entries.  Lossless; a sub-batch with a block whose qpos values do not fit qbits is uploaded raw.
// This kernel restores a[] in HBM.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_unpack_wire(ulonglong2 *__restrict__ a, const uint64_t *__restrict__ packed,
                                                     const uint64_t *__restrict__ tab, const uint32_t *__restrict__ dict, int n)
{
	for (int k = blockIdx.x * 256 + threadIdx.x; k < n; k += gridDim.x * 256) {
		const uint64_t pk = packed[k];
		const uint64_t d = tab[k >> WIRE_BLOCK_SHIFT];
		const unsigned pos = (unsigned)d, nhx = (unsigned)(d >> 32) & 0xffffu, qbits = 24u - (unsigned)(d >> 48);
		const unsigned mid = (unsigned)(pk >> 32) & 0xffffffu;
		const unsigned hx = dict[pos + (unsigned)(pk >> 56)];
		const unsigned hy = dict[pos + nhx + (mid >> qbits)];
		ulonglong2 o;
		o.x = ((uint64_t)hx << 32) | (pk & 0xffffffffull);
		o.y = ((uint64_t)hy << 32) | (mid & ((1u << qbits) - 1u));
		a[k] = o;
	}
}

int launch_unpack_wire(const Workspace &w, const void *d_wire, size_t tab_pos, size_t dict_pos, cudaStream_t s)
{
	const char *base = (const char*)d_wire;
	k_unpack_wire<<<blocks_for(w.n_anchors, 256 * 4, 148 * 32), 256, 0, s>>>((ulonglong2*)w.a, (const uint64_t*)base,
		(const uint64_t*)(base + tab_pos), (const uint32_t*)(base + dict_pos), w.n_anchors);
	return 1;
}

int launch_export_fp(const Workspace &w, cudaStream_t s)
{
	k_export_fp<<<blocks_for(w.n_reads, 8, 148 * 64), 256, 0, s>>>(w);
	return 1;
}

} // namespace mm2b200
