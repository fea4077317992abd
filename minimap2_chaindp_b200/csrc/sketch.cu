// sketch.cu -- mm_sketch (sketch.c:77-143) on the device (include/mm2sketch_b200.h).
//
// The reference walks a read base by base with a ring of the last w k-mers and a running minimum.  The same output follows
// from per-position quantities that need only prefix sums ("sketching"):
//   token   one per base, or one per homopolymer run with HPC (its last base carries the run length); an ambiguous base is a token
//   k-mer   of a valid token = the last k valid tokens (ambiguous bases do not shift the k-mer, they only reset the counters)
//   step    a token that reaches the ring: ambiguous ones (as an empty entry) and valid ones whose k-mer is not its own reverse
//           complement (a symmetric k-mer leaves no trace in l, the ring or its position, sketch.c:109)
//   l       non-symmetric valid tokens since the last ambiguous base; an entry is real iff l >= k and span < 256
//   min     before / after step s = the newest minimum of steps [s-w, s-1] / [s-w+1, s] -- the reference's `min` variable is
//           always exactly that, so what a step pushes (the old minimum when it is replaced or leaves the window, the equal
//           k-mers after a rescan, the first-window special case) is a function of the w+1 entries around it
// One CTA per read walks it in tiles of 512 bases; three block scans per tile place tokens, steps and emitted minimizers.
// A read's minimizers are first written to a slot of len entries (mm_sketch emits at most one per base), then packed.
#include <vector>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include "mm2sketch_b200.h"
#include "seed_internal.h"

using namespace mm2b200;

namespace {

#define FULL 0xffffffffu
#define SK_T 512                   // bases per tile = threads per CTA
#define SK_HALO 256                // ring entries carried between tiles (w < 256)
#define SK_KH 32                   // valid tokens carried between tiles (k <= 28)
#define NONE 0xffffffffffffffffull

struct SketchWork {
	const uint8_t *seq; const int64_t *seq_off; const uint32_t *rid;
	int32_t n_reads, w, k, is_hpc;
	int32_t r_begin, r_end;        // the reads this launch sketches (a batch may arrive in parts)
	ulonglong2 *raw;               // slot of read r at seq_off[r]
	int32_t *n_mv; int64_t *mv_off; ulonglong2 *mv; int64_t *totals;
	int *next_read;                // k_sketch_odd: the launch's work counter (starts at 0)
};

__device__ __forceinline__ uint64_t hash64m(uint64_t key, uint64_t mask)     // sketch.c:27-37
{
	key = (~key + (key << 21)) & mask;
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8)) & mask;
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4)) & mask;
	key = key ^ key >> 28;
	key = (key + (key << 31)) & mask;
	return key;
}

__device__ __forceinline__ int nt4(uint8_t ch)                               // seq_nt4_table, sketch.c:8-25
{
	if (ch < 4) return ch;
	const uint8_t u = ch | 0x20;
	return u == 'a' ? 0 : u == 'c' ? 1 : u == 'g' ? 2 : (u == 't' || u == 'u') ? 3 : 4;
}

// the same function in 32-bit arithmetic: every step is exact modulo 2^(2k), so for 2k <= 32 the low word is all there is
__device__ __forceinline__ uint32_t hash32m(uint32_t key, uint32_t mask)
{
	key = (~key + (key << 21)) & mask;
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8)) & mask;
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4)) & mask;
	key = key ^ key >> 28;
	key = (key + (key << 31)) & mask;
	return key;
}

// inclusive block scan (sum) of a 32-bit value holding up to three 10-bit counters (a tile has 512 positions).  Every scan of
// a tile has its own buf[]: nothing reads it between this call's first write and the previous tile's last read but barriers,
// and buf[SK_T / 32 - 1] keeps the block total until the same scan of the next tile.
__device__ __forceinline__ uint32_t block_scan(uint32_t v, uint32_t *buf, int tid)
{
	const int lane = tid & 31, wid = tid >> 5;
	for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, v, d); if (lane >= d) v += t; }
	if (lane == 31) buf[wid] = v;
	__syncthreads();
	if (wid == 0) {
		uint32_t x = lane < SK_T / 32 ? buf[lane] : 0;
		for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(FULL, x, d); if (lane >= d) x += t; }
		if (lane < SK_T / 32) buf[lane] = x;
	}
	__syncthreads();
	return v + (wid ? buf[wid - 1] : 0);
}

// the 2-bit groups of x in reverse order (group g <-> group 31 - g)
__device__ __forceinline__ uint64_t rev2(uint64_t x)
{
	const uint64_t y = __brevll(x);
	return (y & 0x5555555555555555ull) << 1 | (y >> 1 & 0x5555555555555555ull);
}

__global__ void __launch_bounds__(SK_T) k_sketch(SketchWork W)
{
	__shared__ unsigned long long sx[SK_HALO + SK_T], sy[SK_HALO + SK_T];
	__shared__ int sl[SK_HALO + SK_T];
	__shared__ __align__(16) uint8_t vcode[SK_KH + SK_T];
	__shared__ uint16_t vskip[SK_KH + SK_T];
	__shared__ uint16_t recV[SK_T], recS[SK_T];          // per ambiguous token of the tile: valid / non-symmetric tokens up to it
	__shared__ uint32_t s_wA[SK_T / 32], s_wB[SK_T / 32], s_wE[SK_T / 32];   // one per scan of a tile (block_scan)
	__shared__ uint32_t sk[SK_HALO + SK_T];               // 32-bit image of sx[] >> 8 (the hash) for the fast window pass
	__shared__ uint32_t vpack[(SK_KH + SK_T) / 16 + 2];    // vcode, 16 two-bit codes per word
	__shared__ int c_V0, c_VN0, c_L0, c_E0;                // carried: valid tokens so far (saturated), valid / non-symmetric since the last ambiguous base, emitted
	const int tid = threadIdx.x;
	const int w = W.w, k = W.k, hpc = W.is_hpc;
	const uint64_t mask = (1ULL << 2 * k) - 1;
	const bool fast = !hpc && 2 * k < 32;
	for (int r = W.r_begin + blockIdx.x; r < W.r_end; r += gridDim.x) {
		const int64_t s0 = W.seq_off[r];
		const int len = (int)(W.seq_off[r + 1] - s0);
		const uint8_t *str = W.seq + s0;
		ulonglong2 *out = W.raw + s0;
		const uint32_t rid = W.rid ? W.rid[r] : 0;
		__syncthreads();
		if (tid < SK_HALO) { sx[tid] = NONE; sy[tid] = NONE; sl[tid] = 0; sk[tid] = 0xffffffffu; }
		if (tid < SK_KH) { vcode[tid] = 0; vskip[tid] = 0; }
		if (tid == 0) { c_V0 = 0; c_VN0 = 0; c_L0 = 0; c_E0 = 0; }
		__syncthreads();
		for (int P0 = 0; P0 < len; P0 += SK_T) {
			// ---- tokens
			const int p = P0 + tid;
			const bool active = p < len;
			const int c = active ? nt4(str[p]) : 4;
			bool is_tok = active;
			int skip_len = 1;
			if (hpc && active && c < 4) {
				const int cn = p + 1 < len ? nt4(str[p + 1]) : 5;
				is_tok = cn != c;                            // the last base of a homopolymer run stands for the run
				if (is_tok) while (skip_len < 256 && p - skip_len >= 0 && nt4(str[p - skip_len]) == c) ++skip_len;   // 256+ can never be a real entry (span < 256)
			}
			const bool valid = is_tok && c < 4, isN = is_tok && c >= 4;
			const uint32_t a_sc = block_scan((uint32_t)is_tok | (uint32_t)valid << 10 | (uint32_t)isN << 20, s_wA, tid);
			const int valid_incl = (int)(a_sc >> 10 & 0x3ff), nN_incl = (int)(a_sc >> 20 & 0x3ff);
			const int vr = valid_incl - (valid ? 1 : 0);
			if (valid) { vcode[SK_KH + vr] = (uint8_t)c; vskip[SK_KH + vr] = (uint16_t)skip_len; }
			if (isN) recV[nN_incl - 1] = (uint16_t)valid_incl;
			__syncthreads();
			if (tid < (SK_KH + SK_T) / 16 + 2) {                 // 16 codes -> one word (stale codes behind the tile's last valid token are never read)
				uint32_t word = 0;
				if (tid < (SK_KH + SK_T) / 16) {
					const uint32_t *src = (const uint32_t*)vcode + 4 * tid;
					#pragma unroll
					for (int q = 0; q < 4; ++q) {
						const uint32_t v = src[q];
						word |= ((v & 3) | (v >> 6 & 0xc) | (v >> 12 & 0x30) | (v >> 18 & 0xc0)) << 8 * q;
					}
				}
				vpack[tid] = word;
			}
			__syncthreads();
			// ---- k-mers of the valid tokens
			const int V0 = c_V0, VN0 = c_VN0, L0 = c_L0, E0 = c_E0;
			uint64_t kmer0 = 0, kmer1 = 0;
			int span_hpc = 0;
			bool nonsym = false;
			if (valid) {
				const int before = V0 + vr;                  // valid tokens before this one (saturated, only compared with k <= 28)
				const int q = SK_KH + vr;                    // this token in vcode[]; the 32 codes up to it as one word, oldest lowest
				const int wi = (q - 31) >> 4, sh = 2 * ((q - 31) & 15);
				const uint32_t w0 = vpack[wi], w1 = vpack[wi + 1], w2 = vpack[wi + 2];
				const uint64_t S = (uint64_t)__funnelshift_r(w1, w2, sh) << 32 | __funnelshift_r(w0, w1, sh);
				kmer0 = rev2(S) & mask;                      // newest token lowest; tokens before the read's first one are zero like the reference's initial k-mer
				kmer1 = ~S >> (64 - 2 * k);                  // complement, newest token highest
				if (before < k - 1) kmer1 &= ~((1ULL << 2 * (k - 1 - before)) - 1);     // ... and absent from the reverse k-mer
				if (hpc) {
					const int sinceN = nN_incl > 0 ? valid_incl - (int)recV[nN_incl - 1] : VN0 + valid_incl;
					const int nsum = sinceN < k ? sinceN : k;
					for (int j = 0; j < nsum; ++j) span_hpc += vskip[q - j];
				}
				nonsym = kmer0 != kmer1;
			}
			const bool step = isN || nonsym;
			const uint32_t b_sc = block_scan((uint32_t)nonsym | (uint32_t)step << 10, s_wB, tid);
			const int nonsym_incl = (int)(b_sc & 0x3ff), step_incl = (int)(b_sc >> 10 & 0x3ff);
			if (isN) recS[nN_incl - 1] = (uint16_t)nonsym_incl;
			__syncthreads();
			if (step) {
				uint64_t ix = NONE, iy = NONE;
				int l = 0;
				if (nonsym) {
					l = nN_incl > 0 ? nonsym_incl - (int)recS[nN_incl - 1] : L0 + nonsym_incl;
					const int span = hpc ? span_hpc : (l < k ? l : k);
					if (l >= k && span < 256) {
						const int z = kmer0 < kmer1 ? 0 : 1;
						const uint64_t km = z ? kmer1 : kmer0;
						ix = (k <= 16 ? (uint64_t)hash32m((uint32_t)km, (uint32_t)mask) : hash64m(km, mask)) << 8 | (uint64_t)span;
						iy = (uint64_t)rid << 32 | (uint64_t)((uint32_t)p << 1) | (uint64_t)z;
					}
				}
				const int si = SK_HALO + step_incl - 1;
				sx[si] = ix; sy[si] = iy; sl[si] = l; sk[si] = ix == NONE ? 0xffffffffu : (uint32_t)(ix >> 8);
			}
			__syncthreads();
			// ---- what each step pushes (sketch.c:118-137)
			const uint32_t totA = s_wA[SK_T / 32 - 1], totB = s_wB[SK_T / 32 - 1];      // tile totals, left behind by the scans
			const int tv = (int)(totA >> 10 & 0x3ff), tn = (int)(totA >> 20 & 0x3ff), ts = (int)(totB & 0x3ff), n_steps = (int)(totB >> 10 & 0x3ff);
			int cnt = 0, s = 0, l = 0, prev_i = 0, cur_i = 0;
			uint64_t prev_x = NONE;
			bool emitA = false, emitPrev = false, emitDup = false;
			if (tid < n_steps) {
				s = SK_HALO + tid;
				l = sl[s];
				const uint64_t info_x = sx[s];
				// one pass over the entries both windows share, [s-w+1, s-1]: their newest minimum and how many entries equal it.
				// Without HPC every real entry has span k, so x compares like its hash: a 32-bit pass when the hash leaves
				// 0xffffffff free for the empty entry (2k < 32).
				uint64_t cm = NONE, old_x;
				int ci = s - w, eq = 0;
				if (fast) {
					uint32_t m32 = 0xffffffffu;
					for (int i = s - w + 1; i < s; ++i) {
						const uint32_t x = sk[i];
						if (x < m32) { m32 = x; ci = i; eq = 1; }
						else if (x == m32) { ci = i; ++eq; }
					}
					cm = m32 == 0xffffffffu ? NONE : (uint64_t)m32 << 8 | (uint64_t)k;
					const uint32_t o32 = sk[s - w];
					old_x = o32 == 0xffffffffu ? NONE : (uint64_t)o32 << 8 | (uint64_t)k;
				} else {
					for (int i = s - w + 1; i < s; ++i) {
						const uint64_t x = sx[i];
						if (x < cm) { cm = x; ci = i; eq = 1; }
						else if (x == cm) { ci = i; ++eq; }
					}
					old_x = sx[s - w];
				}
				const bool old_min = old_x < cm;             // the entry that just left wins only when strictly smaller than everything newer
				prev_x = old_min ? old_x : cm;
				prev_i = old_min ? s - w : ci;
				if (l == w + k - 1 && prev_x != NONE && !old_min && eq > 1) { emitA = true; cnt += eq - 1; }
				if (info_x <= prev_x) {                      // the minimum is replaced
					if (l >= w + k && prev_x != NONE) { emitPrev = true; ++cnt; }
				} else if (old_min) {                        // the minimum has left the window: the new one is the newest minimum of [s-w+1, s]
					if (l >= w + k - 1 && prev_x != NONE) { emitPrev = true; ++cnt; }
					const bool take = info_x <= cm;
					cur_i = take ? s : ci;
					const int ndup = take ? (info_x == cm ? eq : 0) : eq - 1;
					if (l >= w + k - 1 && (take ? info_x : cm) != NONE && ndup > 0) { emitDup = true; cnt += ndup; }
				}
			}
			const uint32_t e_sc = block_scan((uint32_t)cnt, s_wE, tid);
			if (cnt) {
				ulonglong2 *o = out + E0 + (int)(e_sc - (uint32_t)cnt);
				if (emitA) for (int i = s - w + 1; i < s; ++i) if (sx[i] == prev_x && i != prev_i) *o++ = make_ulonglong2(sx[i], sy[i]);
				if (emitPrev) *o++ = make_ulonglong2(prev_x, sy[prev_i]);
				if (emitDup) { const uint64_t cur_x = sx[cur_i]; for (int i = s - w + 1; i <= s; ++i) if (sx[i] == cur_x && i != cur_i) *o++ = make_ulonglong2(sx[i], sy[i]); }
			}
			// ---- carry to the next tile: the last SK_HALO ring entries, the last SK_KH valid tokens, the counters
			const int te = (int)s_wE[SK_T / 32 - 1];
			unsigned long long hx = 0, hy = 0; int hl = 0; uint8_t hc = 0; uint16_t hs = 0; uint32_t hk = 0;
			if (tid < SK_HALO) { hx = sx[n_steps + tid]; hy = sy[n_steps + tid]; hl = sl[n_steps + tid]; hk = sk[n_steps + tid]; }
			if (tid < SK_KH) { hc = vcode[tv + tid]; hs = vskip[tv + tid]; }
			__syncthreads();
			if (tid < SK_HALO) { sx[tid] = hx; sy[tid] = hy; sl[tid] = hl; sk[tid] = hk; }
			if (tid < SK_KH) { vcode[tid] = hc; vskip[tid] = hs; }
			if (tid == 0) {
				c_V0 = V0 + tv > 64 ? 64 : V0 + tv;
				c_VN0 = tn > 0 ? tv - (int)recV[tn - 1] : VN0 + tv;
				c_L0 = tn > 0 ? ts - (int)recS[tn - 1] : L0 + ts;
				c_E0 = E0 + te;
			}
			__syncthreads();
		}
		// ---- the minimum that is still current at the end of the read (sketch.c:141-142)
		if (tid == 0) {
			int e = c_E0;
			uint64_t bx = NONE; int bi = SK_HALO - 1;
			for (int i = SK_HALO - w; i < SK_HALO; ++i) if (sx[i] <= bx) { bx = sx[i]; bi = i; }
			if (bx != NONE) out[e++] = make_ulonglong2(bx, sy[bi]);
			W.n_mv[r] = e;
		}
	}
}

// ---- the common presets: no HPC, odd k <= 15 (map-ont, ava-ont: k = 15)
// Three facts make this case much cheaper than the general kernel above:
//  * a k-mer of odd length is never its own reverse complement (its middle base would have to be its own complement; a k-mer
//    that is still filling up cannot be either: its forward image ends in zeros where the reverse image has a complemented
//    base), so EVERY base is a step: the ring index of a base is its position, and the scans of tokens and steps fall away;
//  * l = bases since the last ambiguous one, and an entry is real iff l >= k -- then its k-mer is the k bases up to it, read
//    straight from the 2-bit image of the tile by position (what the k-mer holds while l < k never matters);
//  * x = hash << 8 | k and y = rid << 32 | pos << 1 | strand follow from a 32-bit key and one strand bit per base.
// The ring is a circular buffer of 1024 keys (a tile never looks back further than w < 256) whose first 256 slots are
// mirrored behind its end, so a window is w + 1 consecutive words.  Four block barriers per tile of 512 bases.
#define RK_N 1024
#define RK_EMPTY 0xffffffffu

__device__ __forceinline__ uint32_t rev2_32(uint32_t x)
{
	const uint32_t y = __brev(x);
	return (y & 0x55555555u) << 1 | (y >> 1 & 0x55555555u);
}

template <int TW>                  // w at compile time (0: any w)
__global__ void __launch_bounds__(SK_T, 3) k_sketch_odd(SketchWork W)
{
	__shared__ uint32_t rk[RK_N + SK_HALO];                // keys by position & 1023: the hash, or RK_EMPTY
	__shared__ uint8_t rz[RK_N + SK_HALO];                 // strand of the entry
	__shared__ __align__(8) uint32_t pw[64];               // 2-bit codes of the bases, 16 per word, by (position >> 4) & 63
	__shared__ int s_lastN[SK_T / 32];
	__shared__ uint32_t s_wE[SK_T / 32];
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	const int w = TW ? TW : W.w, k = W.k;
	const uint32_t mask = (1u << 2 * k) - 1;
	__shared__ int s_r;
	for (;;) {
		__syncthreads();
		if (tid == 0) s_r = W.r_begin + atomicAdd(W.next_read, 1);     // reads are handed out one by one: they differ in length
		__syncthreads();
		const int r = s_r;
		if (r >= W.r_end) break;
		const int64_t s0 = W.seq_off[r];
		const int len = (int)(W.seq_off[r + 1] - s0);
		const uint8_t *str = W.seq + s0;
		ulonglong2 *out = W.raw + s0;
		const uint64_t rid_hi = (uint64_t)(W.rid ? W.rid[r] : 0) << 32;
		if (tid < SK_HALO) rk[RK_N - SK_HALO + tid] = RK_EMPTY;       // positions -256 .. -1
		if (tid == 0) pw[63] = 0;
		int carryN = -1, E0 = 0;                           // position of the last ambiguous base so far; minimizers emitted so far
		uint32_t ch_next = tid < len ? str[tid] : 'N';     // the tile's base is fetched one tile ahead
		for (int P0 = 0; P0 < len; P0 += SK_T) {
			const int p = P0 + tid;
			const bool active = p < len;
			// ---- the base: 2-bit code, ambiguity (seq_nt4_table, sketch.c:8-25)
			const uint32_t ch = ch_next;
			ch_next = p + SK_T < len ? str[p + SK_T] : 'N';
			const uint32_t u = (ch | 0x20) - 'a';
			const bool letter = u < 21 && (0x180045u >> u & 1);            // a c g t u
			const bool valid = letter || ch < 4;
			const uint32_t h = ch >> 1 & 3;
			const uint32_t c = letter ? h ^ h >> 1 : ch & 3;               // A C T G -> 0 1 3 2
			const bool isN = active && !valid;
			const uint32_t nb = __ballot_sync(FULL, isN);
			const uint32_t sh = 2 * (lane & 15);
			const uint32_t lo = __reduce_or_sync(FULL, lane < 16 ? c << sh : 0u), hi = __reduce_or_sync(FULL, lane < 16 ? 0u : c << sh);
			if (lane == 0) {
				*(uint2*)&pw[(P0 >> 4 & 63) + 2 * wid] = make_uint2(lo, hi);
				s_lastN[wid] = nb ? P0 + 32 * wid + 31 - __clz(nb) : -1;
			}
			const int anyN = __syncthreads_or(nb != 0);
			int lastN = carryN;
			if (anyN) {                                    // rare: a tile with ambiguous bases
				const uint32_t mine = nb & (0xffffffffu >> (31 - lane));
				int t = carryN;
				for (int j = 0; j < SK_T / 32; ++j) { const int v = s_lastN[j]; if (j < wid && v > lastN) lastN = v; if (v > t) t = v; }
				if (mine) lastN = P0 + 32 * wid + 31 - __clz(mine);
				carryN = t;
			}
			// ---- k-mer, key, strand
			uint32_t key = RK_EMPTY, z = 0;
			const int l = valid ? p - lastN : 0;           // isN: lastN == p
			if (active && l >= k) {
				const int q = p - 15, a = q >> 4;
				const uint32_t V = __funnelshift_r(pw[a & 63], pw[(a + 1) & 63], 2 * (q & 15));   // the 16 bases up to p, oldest lowest
				const uint32_t kmer0 = rev2_32(V) & mask, kmer1 = ~V >> (32 - 2 * k);
				z = kmer0 < kmer1 ? 0 : 1;
				key = hash32m(z ? kmer1 : kmer0, mask);
			}
			const int slot = p & (RK_N - 1);
			if (active) {
				rk[slot] = key; rz[slot] = (uint8_t)z;
				if (slot < SK_HALO) { rk[RK_N + slot] = key; rz[RK_N + slot] = (uint8_t)z; }
			}
			__syncthreads();
			// ---- what the step pushes (sketch.c:118-137); entry j of the window is position p - w + j
			const int base = (p - w) & (RK_N - 1);
			const uint32_t *win = rk + base;
			int cnt = 0, prev_j = 0, cur_j = 0;
			uint32_t prev = RK_EMPTY, m = RK_EMPTY;
			bool emitA = false, emitPrev = false, emitDup = false;
			if (active) {
				int cj = 0, eq = 0;                        // newest entry that equals the minimum, and how many do
				if (TW > 0 && TW <= 32) {                  // unrolled: the minimum first, then a bit mask of the entries equal to it
					uint32_t x[TW > 0 ? TW : 1], em = 0;
					#pragma unroll
					for (int j = 1; j < TW; ++j) { x[j] = win[j]; m = x[j] < m ? x[j] : m; }
					#pragma unroll
					for (int j = 1; j < TW; ++j) if (x[j] == m) em |= 1u << j;
					eq = __popc(em);
					cj = em ? 31 - __clz(em) : 0;
				} else {
					for (int j = 1; j < w; ++j) {
						const uint32_t x = win[j];
						if (x < m) { m = x; cj = j; eq = 1; }
						else if (x == m) { cj = j; ++eq; }
					}
				}
				const uint32_t old = win[0];
				const bool old_min = old < m;
				prev = old_min ? old : m;
				prev_j = old_min ? 0 : cj;
				if (l == w + k - 1 && prev != RK_EMPTY && !old_min && eq > 1) { emitA = true; cnt += eq - 1; }
				if (key <= prev) {
					if (l >= w + k && prev != RK_EMPTY) { emitPrev = true; ++cnt; }
				} else if (old_min) {
					if (l >= w + k - 1 && prev != RK_EMPTY) { emitPrev = true; ++cnt; }
					const bool take = key <= m;
					cur_j = take ? w : cj;
					const int ndup = take ? (key == m ? eq : 0) : eq - 1;
					if (l >= w + k - 1 && (take ? key : m) != RK_EMPTY && ndup > 0) { emitDup = true; cnt += ndup; }
				}
			}
			const uint32_t e_sc = block_scan((uint32_t)cnt, s_wE, tid);
			if (cnt) {
				ulonglong2 *o = out + E0 + (int)(e_sc - (uint32_t)cnt);
				const uint8_t *wz = rz + base;
				auto y_of = [&](int j) { return rid_hi | (uint64_t)((uint32_t)(p - w + j) << 1) | (uint64_t)wz[j]; };
				if (emitA) for (int j = 1; j < w; ++j) if (win[j] == prev && j != prev_j) *o++ = make_ulonglong2((uint64_t)prev << 8 | (uint64_t)k, y_of(j));
				if (emitPrev) *o++ = make_ulonglong2((uint64_t)prev << 8 | (uint64_t)k, y_of(prev_j));
				if (emitDup) {
					const uint32_t cur = cur_j == w ? key : m;
					for (int j = 1; j <= w; ++j) if (win[j] == cur && j != cur_j) *o++ = make_ulonglong2((uint64_t)cur << 8 | (uint64_t)k, y_of(j));
				}
			}
			E0 += (int)s_wE[SK_T / 32 - 1];
		}
		// ---- the minimum that is still current at the end of the read (sketch.c:141-142)
		if (tid == 0) {
			uint32_t bx = RK_EMPTY; int bi = 0;
			for (int i = len - w; i < len; ++i) { const uint32_t x = rk[i & (RK_N - 1)]; if (x <= bx) { bx = x; bi = i; } }
			if (bx != RK_EMPTY && len > 0) out[E0++] = make_ulonglong2((uint64_t)bx << 8 | (uint64_t)k, rid_hi | (uint64_t)((uint32_t)bi << 1) | rz[bi & (RK_N - 1)]);
			W.n_mv[r] = E0;
		}
	}
}

// exclusive scan of n_mv over the reads by one CTA
__global__ void __launch_bounds__(1024) k_sketch_scan(SketchWork W)
{
	__shared__ long long s_warp[32];
	__shared__ long long s_carry;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	if (tid == 0) s_carry = 0;
	__syncthreads();
	for (int r0 = 0; r0 < W.n_reads; r0 += 1024) {
		const int r = r0 + tid;
		const long long v = r < W.n_reads ? W.n_mv[r] : 0;
		long long incl = v;
		for (int d = 1; d < 32; d <<= 1) { const long long t = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += t; }
		if (lane == 31) s_warp[wid] = incl;
		__syncthreads();
		if (wid == 0) {
			long long x = s_warp[lane];
			for (int d = 1; d < 32; d <<= 1) { const long long t = __shfl_up_sync(FULL, x, d); if (lane >= d) x += t; }
			s_warp[lane] = x;
		}
		__syncthreads();
		const long long before = s_carry + (wid ? s_warp[wid - 1] : 0) + incl - v;
		if (r < W.n_reads) W.mv_off[r] = before;
		__syncthreads();
		if (tid == 1023) s_carry = before + v;
		__syncthreads();
	}
	if (tid == 0) { W.mv_off[W.n_reads] = s_carry; W.totals[0] = s_carry; }
}

__global__ void __launch_bounds__(256) k_sketch_pack(SketchWork W)
{
	for (int r = blockIdx.x; r < W.n_reads; r += gridDim.x) {
		const int n = W.n_mv[r];
		const ulonglong2 *src = W.raw + W.seq_off[r];
		ulonglong2 *dst = W.mv + W.mv_off[r];
		for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
	}
}

inline bool getenv_off(const char *name) { const char *e = getenv(name); return e && e[0] == '0'; }
inline size_t up256(size_t x) { return (x + 255) / 256 * 256; }

struct Buf {
	void *p = nullptr; size_t cap = 0; bool pinned = false;
	int reserve(size_t bytes);
	void release() { if (p) { if (pinned) cudaFreeHost(p); else cudaFree(p); } p = nullptr; cap = 0; }
};

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

int Buf::reserve(size_t bytes)
{
	if (p && bytes <= cap) return MM2B200_OK;
	release();
	const size_t want = bytes + bytes / 8 + 4096;
	const cudaError_t e = pinned ? cudaHostAlloc(&p, want, cudaHostAllocDefault) : cudaMalloc(&p, want);
	if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; return failf(MM2B200_ERR_NOMEM, "%s(%zu bytes) failed in the sketching stage", pinned ? "cudaHostAlloc" : "cudaMalloc", want); }
	cap = want;
	return MM2B200_OK;
}

struct SketchState {
	int device = 0;
	std::mutex mu;
	cudaStream_t stream = nullptr;
	Buf d_raw, d_mv, d_small, d_in;
	Buf h_off, h_mv, h_tot;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	// pipelined form (sketch_pipe_launch / _finish): where the group's arrays lie in d_raw / d_small
	ulonglong2 *p_raw = nullptr; int32_t *p_n_mv = nullptr; int64_t *p_mv_off = nullptr, *p_totals = nullptr; int *p_counters = nullptr;
	bool pipe_timed = false;
	double ms = 0; int64_t launches = 0;
	SketchState() { h_off.pinned = h_mv.pinned = h_tot.pinned = true; }
};

void sketch_state_free(void *p)
{
	SketchState *s = (SketchState*)p;
	cudaSetDevice(s->device);
	if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
	for (Buf *b : { &s->d_raw, &s->d_mv, &s->d_small, &s->d_in, &s->h_off, &s->h_mv, &s->h_tot }) b->release();
	if (s->ev0) cudaEventDestroy(s->ev0);
	if (s->ev1) cudaEventDestroy(s->ev1);
	delete s;
}

std::mutex g_state_mu;
SketchState *sketch_state(mm2b200_ctx_t *ctx)
{
	std::lock_guard<std::mutex> lk(g_state_mu);
	SketchState *s = (SketchState*)ctx_ext(ctx, 1);
	if (!s) {
		s = new SketchState();
		s->device = ctx_device(ctx);
		ctx_set_ext(ctx, 1, s, sketch_state_free);
	}
	return s;
}

int check_sketch(const mm2b200_sketch_params_t *prm)
{
	if (!prm) return failf(MM2B200_ERR_ARG, "sketching parameters are NULL");
	if (prm->w <= 0 || prm->w >= 256 || prm->k <= 0 || prm->k > 28)            // the assert of sketch.c:84
		return failf(MM2B200_ERR_ARG, "mm_sketch needs 0 < w < 256 and 0 < k <= 28 (w %d, k %d)", prm->w, prm->k);
	return MM2B200_OK;
}

// kernels on `st`; waits once for the count; leaves the packing on the stream
int sketch_device_impl(mm2b200_ctx_t *ctx, SketchState *s, const mm2b200_sketch_params_t *prm, int64_t n_reads, int64_t n_bases,
                       const int64_t *d_seq_off, const char *d_seq, const uint32_t *d_rid, cudaStream_t st, mm2b200_dev_sketch_t *out, int64_t *n_mv)
{
	int rc;
	const size_t R = (size_t)n_reads;
	if ((rc = s->d_raw.reserve(16 * ((size_t)n_bases + 1))) || (rc = s->d_small.reserve(up256(4 * (R + 1)) + up256(8 * (R + 2)) + 256 + 256)) ||
	    (rc = s->h_tot.reserve(64))) return rc;
	SketchWork W;
	memset(&W, 0, sizeof(W));
	W.seq = (const uint8_t*)d_seq; W.seq_off = d_seq_off; W.rid = d_rid;
	W.n_reads = (int32_t)n_reads; W.w = prm->w; W.k = prm->k; W.is_hpc = prm->is_hpc ? 1 : 0;
	W.raw = (ulonglong2*)s->d_raw.p;
	char *cur = (char*)s->d_small.p;
	W.n_mv = (int32_t*)cur; cur += up256(4 * (R + 1));
	W.mv_off = (int64_t*)cur; cur += up256(8 * (R + 2));
	W.totals = (int64_t*)cur; cur += 256;
	int *counters = (int*)cur;                                         // the work counter of k_sketch_odd
	const bool prof = ctx_profile(ctx) != 0;
	if (prof) {
		if (!s->ev0) { CU_TRY(cudaEventCreate(&s->ev0)); CU_TRY(cudaEventCreate(&s->ev1)); }
		CU_TRY(cudaEventRecord(s->ev0, st));
	}
	int64_t total = 0;
	int launches = 0;
	if (n_reads > 0) {
		const int sm = ctx_sm_count(ctx);
		const int64_t resident = (int64_t)sm * 3;                       // both kernels: 512 threads of 40 registers
		int grid = (int)(n_reads < resident ? n_reads : resident);
		// no HPC and an odd k <= 15: the specialised kernel (w = 5 and 10 unrolled: ava-ont, map-ont)
		const bool odd = !W.is_hpc && (W.k & 1) && W.k <= 15 && !getenv_off("MM2B200_SKETCH_ODD");
		auto launch = [&](int g) {
			if (!odd) k_sketch<<<g, SK_T, 0, st>>>(W);
			else if (W.w == 10) k_sketch_odd<10><<<g, SK_T, 0, st>>>(W);
			else if (W.w == 5) k_sketch_odd<5><<<g, SK_T, 0, st>>>(W);
			else k_sketch_odd<0><<<g, SK_T, 0, st>>>(W);
		};
		if (odd) CU_TRY(cudaMemsetAsync(counters, 0, 4, st));
		W.next_read = counters;
		W.r_begin = 0; W.r_end = (int32_t)n_reads;
		launch(grid);
		k_sketch_scan<<<1, 1024, 0, st>>>(W);
		CU_TRY(cudaGetLastError());
		CU_TRY(cudaMemcpyAsync(s->h_tot.p, W.totals, 8, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaStreamSynchronize(st));
		total = *(int64_t*)s->h_tot.p;
		if ((rc = s->d_mv.reserve(16 * ((size_t)total + 1)))) return rc;
		W.mv = (ulonglong2*)s->d_mv.p;
		k_sketch_pack<<<(int)(n_reads < (int64_t)sm * 8 ? n_reads : (int64_t)sm * 8), 256, 0, st>>>(W);
		CU_TRY(cudaGetLastError());
		launches += 3;
	} else {
		if ((rc = s->d_mv.reserve(16))) return rc;
		CU_TRY(cudaMemsetAsync(W.mv_off, 0, 8, st));
		CU_TRY(cudaMemsetAsync(W.totals, 0, 64, st));
	}
	if (prof) CU_TRY(cudaEventRecord(s->ev1, st));
	s->launches += launches;
	ctx_add_launches(ctx, launches);
	out->d_mv_off = W.mv_off; out->d_mv = (const mm128_t*)s->d_mv.p; out->d_totals = W.totals;
	if (n_mv) *n_mv = total;
	return MM2B200_OK;
}

void sketch_account(mm2b200_ctx_t *ctx, SketchState *s)
{
	if (ctx_profile(ctx) && s->ev0) {
		float t = 0;
		if (cudaEventElapsedTime(&t, s->ev0, s->ev1) == cudaSuccess) s->ms += t;
		cudaGetLastError();
	}
}

} // namespace

// The composed path, pipelined (seed_api.cu): a GROUP of chunks shares one slot array and one count array.  _launch puts the
// sketching kernels of one chunk's parts on the sketching stream (they follow the upload) and records done_ev behind them;
// _finish, on the compute stream, waits for done_ev, scans the chunk's counts and packs its minimizers -- so the seeding and
// chaining of a chunk run while the next chunk's bases are still being uploaded and sketched.  r_off / b_off: the chunk's
// first read / base inside the group; d_seq_off are the chunk's own offsets (first = 0).  `first` sizes the group's arrays
// (the caller has synchronised everything that used them).
namespace mm2b200 {
int sketch_pipe_launch(mm2b200_ctx *ctx, const mm2b200_sketch_params_t *prm, bool first, bool last, int64_t group_reads, int64_t group_bases, int group_parts,
                       int64_t r_off, int64_t b_off, int part_off, int64_t nr, const int64_t *d_seq_off, const char *d_seq, cudaStream_t st,
                       int n_parts, const int64_t *part_cut, const cudaEvent_t *part_ev, cudaEvent_t done_ev)
{
	int rc;
	if ((rc = check_sketch(prm))) return rc;
	SketchState *s = sketch_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	CU_TRY(cudaSetDevice(s->device));
	const bool prof = ctx_profile(ctx) != 0;
	if (first) {
		const size_t R = (size_t)group_reads;
		if ((rc = s->d_raw.reserve(16 * ((size_t)group_bases + 1))) ||
		    (rc = s->d_small.reserve(up256(4 * (R + 1)) + up256(8 * (R + 2)) + 256 + up256(4 * ((size_t)group_parts + 1)))) || (rc = s->h_tot.reserve(64))) return rc;
		char *cur = (char*)s->d_small.p;
		s->p_raw = (ulonglong2*)s->d_raw.p;
		s->p_n_mv = (int32_t*)cur; cur += up256(4 * (R + 1));
		s->p_mv_off = (int64_t*)cur; cur += up256(8 * (R + 2));
		s->p_totals = (int64_t*)cur; cur += 256;
		s->p_counters = (int*)cur;
		CU_TRY(cudaMemsetAsync(s->p_counters, 0, 4 * ((size_t)group_parts + 1), st));
		s->pipe_timed = false;
		if (prof) {
			if (!s->ev0) { CU_TRY(cudaEventCreate(&s->ev0)); CU_TRY(cudaEventCreate(&s->ev1)); }
			CU_TRY(cudaEventRecord(s->ev0, st));
		}
	}
	SketchWork W;
	memset(&W, 0, sizeof(W));
	W.seq = (const uint8_t*)d_seq; W.seq_off = d_seq_off; W.rid = nullptr;
	W.n_reads = (int32_t)nr; W.w = prm->w; W.k = prm->k; W.is_hpc = prm->is_hpc ? 1 : 0;
	W.raw = s->p_raw + b_off; W.n_mv = s->p_n_mv + r_off;
	const int sm = ctx_sm_count(ctx);
	const int64_t resident = (int64_t)sm * 3;
	const bool odd = !W.is_hpc && (W.k & 1) && W.k <= 15 && !getenv_off("MM2B200_SKETCH_ODD");
	int launches = 0;
	for (int k = 0; k < n_parts; ++k) {
		const int64_t n = part_cut[k + 1] - part_cut[k];
		CU_TRY(cudaStreamWaitEvent(st, part_ev[k], 0));
		if (n <= 0) continue;
		W.r_begin = (int32_t)part_cut[k]; W.r_end = (int32_t)part_cut[k + 1]; W.next_read = s->p_counters + part_off + k;
		const int g = (int)(n < resident ? n : resident);
		if (!odd) k_sketch<<<g, SK_T, 0, st>>>(W);
		else if (W.w == 10) k_sketch_odd<10><<<g, SK_T, 0, st>>>(W);
		else if (W.w == 5) k_sketch_odd<5><<<g, SK_T, 0, st>>>(W);
		else k_sketch_odd<0><<<g, SK_T, 0, st>>>(W);
		++launches;
	}
	CU_TRY(cudaGetLastError());
	CU_TRY(cudaEventRecord(done_ev, st));
	if (last && prof) { CU_TRY(cudaEventRecord(s->ev1, st)); s->pipe_timed = true; }
	s->launches += launches;
	ctx_add_launches(ctx, launches);
	return MM2B200_OK;
}

int sketch_pipe_finish(mm2b200_ctx *ctx, int64_t r_off, int64_t b_off, int64_t nr, const int64_t *d_seq_off, cudaStream_t st, cudaEvent_t done_ev,
                       mm2b200_dev_sketch_t *out, int64_t *n_mv)
{
	int rc;
	SketchState *s = sketch_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	CU_TRY(cudaSetDevice(s->device));
	SketchWork W;
	memset(&W, 0, sizeof(W));
	W.seq_off = d_seq_off; W.n_reads = (int32_t)nr;
	W.raw = s->p_raw + b_off; W.n_mv = s->p_n_mv + r_off; W.mv_off = s->p_mv_off; W.totals = s->p_totals;
	CU_TRY(cudaStreamWaitEvent(st, done_ev, 0));
	int64_t total = 0;
	if (nr > 0) {
		const int sm = ctx_sm_count(ctx);
		k_sketch_scan<<<1, 1024, 0, st>>>(W);
		CU_TRY(cudaGetLastError());
		CU_TRY(cudaMemcpyAsync(s->h_tot.p, W.totals, 8, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaStreamSynchronize(st));
		total = *(int64_t*)s->h_tot.p;
		if ((rc = s->d_mv.reserve(16 * ((size_t)total + 1)))) return rc;
		W.mv = (ulonglong2*)s->d_mv.p;
		k_sketch_pack<<<(int)(nr < (int64_t)sm * 8 ? nr : (int64_t)sm * 8), 256, 0, st>>>(W);
		CU_TRY(cudaGetLastError());
		s->launches += 2;
		ctx_add_launches(ctx, 2);
	} else {
		if ((rc = s->d_mv.reserve(16))) return rc;
		CU_TRY(cudaMemsetAsync(W.mv_off, 0, 8, st));
		CU_TRY(cudaMemsetAsync(W.totals, 0, 64, st));
	}
	out->d_mv_off = W.mv_off; out->d_mv = (const mm128_t*)s->d_mv.p; out->d_totals = W.totals;
	if (n_mv) *n_mv = total;
	return MM2B200_OK;
}
}

// after the caller has synchronised the stream: add the events' span to the stage time (the composed path, seed_api.cu)
namespace mm2b200 { void sketch_stage_account(mm2b200_ctx *ctx) { SketchState *s = sketch_state(ctx); std::lock_guard<std::mutex> lk(s->mu); sketch_account(ctx, s); } }

extern "C" int mm2b200_sketch_device(mm2b200_ctx_t *ctx, const mm2b200_sketch_params_t *prm, int64_t n_reads, int64_t n_bases, const int64_t *d_seq_off,
                                     const char *d_seq, const uint32_t *d_rid, void *stream, mm2b200_dev_sketch_t *out, int64_t *n_mv)
{
	int rc;
	if (!ctx || !out) return failf(MM2B200_ERR_ARG, "NULL ctx or result");
	if ((rc = check_sketch(prm))) return rc;
	if (n_reads < 0 || n_bases < 0 || !d_seq_off || (n_bases && !d_seq)) return failf(MM2B200_ERR_ARG, "bad sketching arguments");
	SketchState *s = sketch_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	CU_TRY(cudaSetDevice(s->device));
	return sketch_device_impl(ctx, s, prm, n_reads, n_bases, d_seq_off, d_seq, d_rid, (cudaStream_t)stream, out, n_mv);
}

extern "C" int mm2b200_sketch_stage_ms(mm2b200_ctx_t *ctx, double *ms, int64_t *launches)
{
	if (!ctx) return failf(MM2B200_ERR_ARG, "NULL context");
	SketchState *s = sketch_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	if (ms) *ms = s->ms;
	if (launches) *launches = s->launches;
	return MM2B200_OK;
}

extern "C" int mm2b200_sketch_batch(mm2b200_ctx_t *ctx, const mm2b200_sketch_params_t *prm, int64_t n_reads, const int64_t *seq_off,
                                    const char *seq, const uint32_t *rid, mm2b200_sketch_result_t *res)
{
	int rc;
	if (!ctx || !res) return failf(MM2B200_ERR_ARG, "NULL ctx or result");
	if ((rc = check_sketch(prm))) return rc;
	if (n_reads < 0 || (n_reads && !seq_off)) return failf(MM2B200_ERR_ARG, "bad batch arguments");
	for (int64_t r = 0; r < n_reads; ++r) {
		if (seq_off[r + 1] < seq_off[r]) return failf(MM2B200_ERR_ARG, "seq_off is not non-decreasing at read %lld", (long long)r);
		if (seq_off[r + 1] - seq_off[r] >= ((int64_t)1 << 31) - 1) return failf(MM2B200_ERR_ARG, "read %lld is longer than mm_sketch's int length", (long long)r);
	}
	const int64_t b0 = n_reads ? seq_off[0] : 0, B = n_reads ? seq_off[n_reads] - b0 : 0;
	if (B && !seq) return failf(MM2B200_ERR_ARG, "NULL sequence array");
	SketchState *s = sketch_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	memset(res, 0, sizeof(*res));
	CU_TRY(cudaSetDevice(s->device));
	if (!s->stream) CU_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
	cudaStream_t st = s->stream;
	const size_t R = (size_t)n_reads;
	if ((rc = s->d_in.reserve(up256((size_t)B + 16) + up256(8 * (R + 1)) + up256(4 * (R + 1)))) || (rc = s->h_off.reserve(16 * (R + 1) + 64))) return rc;
	char *cur = (char*)s->d_in.p;
	char *d_seq = cur; cur += up256((size_t)B + 16);
	int64_t *d_off = (int64_t*)cur; cur += up256(8 * (R + 1));
	uint32_t *d_rid = rid ? (uint32_t*)cur : nullptr;
	int64_t *h_off = (int64_t*)s->h_off.p;
	for (size_t r = 0; r <= R; ++r) h_off[r] = n_reads ? seq_off[r] - b0 : 0;
	CU_TRY(cudaMemcpyAsync(d_off, h_off, 8 * (R + 1), cudaMemcpyHostToDevice, st));
	if (B) CU_TRY(cudaMemcpyAsync(d_seq, seq + b0, (size_t)B, cudaMemcpyHostToDevice, st));
	if (rid && R) CU_TRY(cudaMemcpyAsync(d_rid, rid, 4 * R, cudaMemcpyHostToDevice, st));
	mm2b200_dev_sketch_t dev;
	int64_t total = 0;
	if ((rc = sketch_device_impl(ctx, s, prm, n_reads, B, d_off, d_seq, d_rid, st, &dev, &total))) return rc;
	if ((rc = s->h_mv.reserve(16 * ((size_t)total + 1)))) return rc;
	int64_t *h_mv_off = h_off + (R + 1);
	CU_TRY(cudaMemcpyAsync(h_mv_off, dev.d_mv_off, 8 * (R + 1), cudaMemcpyDeviceToHost, st));
	if (total) CU_TRY(cudaMemcpyAsync(s->h_mv.p, dev.d_mv, 16 * (size_t)total, cudaMemcpyDeviceToHost, st));
	CU_TRY(cudaStreamSynchronize(st));
	sketch_account(ctx, s);
	res->n_reads = n_reads; res->mv_off = h_mv_off; res->mv = (const mm128_t*)s->h_mv.p;
	return MM2B200_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// mm2b200_sketch_device_noalloc (include/mm2sketch_b200.h): the sketching kernels with their scratch in a static device
// array, for callers that must not allocate.  No counterpart in the reference.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

// scratch of one call: the minimizer slots (16 B per base), the per-read counts, the work counter of k_sketch_odd.
// Held by every GPU that loads the library until the process exits.
#define NA_RAW_BYTES   ((size_t)16 * (MM2B200_NOALLOC_MAX_BASES + 1))
#define NA_CNT_BYTES   ((size_t)4 * (MM2B200_NOALLOC_MAX_READS + 64))
#define NA_BYTES       (NA_RAW_BYTES + NA_CNT_BYTES + 256)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// In host code g_noalloc_scratch is only a handle: the device address is asked for once per device (that device current).
std::mutex g_na_mu;
unsigned char *g_na_ptr[64];
int g_na_sm[64];

int noalloc_scratch(unsigned char **ptr, int *sm)
{
	int dev = -1;
	CU_TRY(cudaGetDevice(&dev));
	if (dev < 0 || dev >= 64) return failf(MM2B200_ERR_ARG, "device ordinal %d is outside the table of the allocation-free entry point", dev);
	std::lock_guard<std::mutex> lk(g_na_mu);
	if (!g_na_ptr[dev]) {
		unsigned char *p = nullptr;
		CU_TRY(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
		int n = 0;
		CU_TRY(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
		g_na_sm[dev] = n > 0 ? n : 1;
		g_na_ptr[dev] = p;
	}
	*ptr = g_na_ptr[dev]; *sm = g_na_sm[dev];
	return MM2B200_OK;
}

// k_sketch_pack with a bound: the caller's output holds mv_cap minimizers; totals[1] = 1 when the batch has more
__global__ void __launch_bounds__(256) k_sketch_pack_cap(SketchWork W, int64_t mv_cap)
{
	if (W.mv_off[W.n_reads] > mv_cap) {
		if (blockIdx.x == 0 && threadIdx.x == 0) W.totals[1] = 1;
		return;
	}
	for (int r = blockIdx.x; r < W.n_reads; r += gridDim.x) {
		const int n = W.n_mv[r];
		const ulonglong2 *src = W.raw + W.seq_off[r];
		ulonglong2 *dst = W.mv + W.mv_off[r];
		for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
	}
}

} // namespace

extern "C" int mm2b200_sketch_device_noalloc(const mm2b200_sketch_params_t *prm, int64_t n_reads, int64_t n_bases, const int64_t *d_seq_off,
                                             const char *d_seq, const uint32_t *d_rid, void *stream, mm128_t *d_mv, int64_t mv_cap,
                                             int64_t *d_mv_off, int64_t *d_totals, int64_t *n_mv)
{
	int rc;
	if ((rc = check_sketch(prm))) return rc;
	if (n_reads < 0 || n_bases < 0 || mv_cap < 0 || !d_seq_off || !d_mv_off || !d_totals || (n_bases && !d_seq) || (mv_cap && !d_mv))
		return failf(MM2B200_ERR_ARG, "bad sketching arguments");
	if (n_reads > MM2B200_NOALLOC_MAX_READS || n_bases > MM2B200_NOALLOC_MAX_BASES)
		return failf(MM2B200_ERR_ARG, "batch of %lld reads / %lld bases exceeds the static scratch of the allocation-free entry point (%d reads, %d bases): "
		             "send less per call or use mm2b200_sketch_device", (long long)n_reads, (long long)n_bases, MM2B200_NOALLOC_MAX_READS, MM2B200_NOALLOC_MAX_BASES);
	unsigned char *scratch = nullptr;
	int sm = 1;
	if ((rc = noalloc_scratch(&scratch, &sm))) return rc;
	cudaStream_t st = (cudaStream_t)stream;
	SketchWork W;
	memset(&W, 0, sizeof(W));
	W.seq = (const uint8_t*)d_seq; W.seq_off = d_seq_off; W.rid = d_rid;
	W.n_reads = (int32_t)n_reads; W.w = prm->w; W.k = prm->k; W.is_hpc = prm->is_hpc ? 1 : 0;
	W.r_begin = 0; W.r_end = (int32_t)n_reads;
	W.raw = (ulonglong2*)scratch;
	W.n_mv = (int32_t*)(scratch + NA_RAW_BYTES);
	W.next_read = (int*)(scratch + NA_RAW_BYTES + NA_CNT_BYTES);
	W.mv_off = d_mv_off; W.mv = (ulonglong2*)d_mv; W.totals = d_totals;
	// the array keeps what the last call left: the work counter and the totals must start at zero (the slots and counts are
	// written before they are read)
	CU_TRY(cudaMemsetAsync(W.next_read, 0, 256, st));
	CU_TRY(cudaMemsetAsync(d_totals, 0, 64, st));
	if (n_reads > 0) {
		const int64_t resident = (int64_t)sm * 3;
		const int grid = (int)(n_reads < resident ? n_reads : resident);
		const bool odd = !W.is_hpc && (W.k & 1) && W.k <= 15;
		if (!odd) k_sketch<<<grid, SK_T, 0, st>>>(W);
		else if (W.w == 10) k_sketch_odd<10><<<grid, SK_T, 0, st>>>(W);
		else if (W.w == 5) k_sketch_odd<5><<<grid, SK_T, 0, st>>>(W);
		else k_sketch_odd<0><<<grid, SK_T, 0, st>>>(W);
		k_sketch_scan<<<1, 1024, 0, st>>>(W);
		k_sketch_pack_cap<<<(int)(n_reads < (int64_t)sm * 8 ? n_reads : (int64_t)sm * 8), 256, 0, st>>>(W, mv_cap);
		CU_TRY(cudaGetLastError());
	} else CU_TRY(cudaMemsetAsync(d_mv_off, 0, 8, st));
	if (n_mv) {                                   // the caller wants the count now: one wait, and the overflow check with it
		int64_t tot[2] = { 0, 0 };
		CU_TRY(cudaMemcpyAsync(tot, d_totals, 16, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaStreamSynchronize(st));
		*n_mv = tot[0];
		if (tot[1]) return failf(MM2B200_ERR_ARG, "the batch has %lld minimizers, the output holds %lld (nothing was written to it)", (long long)tot[0], (long long)mv_cap);
	}
	return MM2B200_OK;
}
