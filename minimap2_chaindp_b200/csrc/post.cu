// post.cu -- chain -> region post-processing on the device (include/mm2post_b200.h):
// mm_gen_regs (hit.c:52-88) and chain_post (map.c:238-247: mm_set_parent hit.c:109-165, mm_select_sub hit.c:230-247,
// mm_join_long hit.c:290-345 with mm_squeeze_a / mm_filter_regs / mm_sync_regs) on the chains where the DP left them.
//
// One warp per read.  What is data-parallel runs across the lanes: the hashes and offsets of the chains, the region records,
// the coordinate / fuzzy-length pass over every chain anchor (hit.c:8-38: the only O(anchors) part) and the anchor moves of
// mm_squeeze_a.  The order-dependent bookkeeping (who is whose parent, which secondaries survive, which neighbours join) is
// executed by all 32 lanes in lock step on the same values -- same addresses, one broadcast transaction per load, no
// divergence -- with lane 0 doing the stores; it is O(regions^2) on a handful of regions per read.  The sorts that decide an
// ORDER (radix_sort_128x in mm_gen_regs) replay the reference's unstable sort exactly (sort_exact.cuh); the key-only sorts
// (radix_sort_64 on distinct or payload-free keys) use the same routine as a plain correct sort.
// The float expressions are evaluated with explicit round-to-nearest intrinsics in the reference's operation order.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <vector>
#include "mm2post_b200.h"
#include "seed_internal.h"
#include "sort_exact.cuh"

using namespace mm2b200;

namespace {

#define FULL 0xffffffffu
#define PARENT_UNSET   (-1)          // mmpriv.h:8
#define PARENT_TMP_PRI (-2)          // mmpriv.h:9
#define SEED_LONG_JOIN (1ULL << 40)  // mmpriv.h:16
#define F_SPLICE 0x080
#define F_NO_LJOIN 0x400
#define F_SR 0x1000
#define F_ALL_CHAINS 0x800000

typedef mm2b200_reg_t Reg;
static_assert(sizeof(Reg) == 80, "mm_reg1_t is 80 bytes (minimap.h:100-115)");

struct PostWork {
	// chains of the batch as mm2b200_chain_device left them
	const int32_t *n_u; const int64_t *u_off, *b_off; const uint64_t *u; ulonglong2 *b;
	const uint32_t *hash; const int32_t *qlen;
	int32_t n_reads;
	// scratch, indexed by chain (read r owns [u_off[r], u_off[r+1]))
	uint64_t *zx, *zy, *cov, *aux;
	int32_t *w, *tmp, *stk;
	Reg *R;
	// outputs
	int32_t *n_regs; int64_t *reg_off; Reg *out; int64_t *totals;
	mm2b200_post_params_t prm;
};

__device__ __forceinline__ uint64_t hash64(uint64_t key)          // hit.c:40-50
{
	key = (~key + (key << 21));
	key = key ^ key >> 24;
	key = ((key + (key << 3)) + (key << 8));
	key = key ^ key >> 14;
	key = ((key + (key << 2)) + (key << 4));
	key = key ^ key >> 28;
	key = (key + (key << 31));
	return key;
}

__device__ __forceinline__ ulonglong2 ld_anchor(const ulonglong2 *p)   // re-read from memory (lane 0 may just have rewritten it)
{
	ulonglong2 v;
	v.x = ((const volatile unsigned long long*)p)[0];
	v.y = ((const volatile unsigned long long*)p)[1];
	return v;
}

__device__ __forceinline__ void copy_reg(Reg *dst, const Reg *src)   // 80 bytes, 16-byte aligned
{
	const uint4 *s = (const uint4*)src; uint4 *d = (uint4*)dst;
	uint4 v0 = s[0], v1 = s[1], v2 = s[2], v3 = s[3], v4 = s[4];
	d[0] = v0; d[1] = v1; d[2] = v2; d[3] = v3; d[4] = v4;
}

// mm_reg_set_coor + mm_cal_fuzzy_len (hit.c:8-38) by the whole warp: lanes stride over the region's anchors
__device__ void set_coor_warp(volatile Reg *r, int32_t qlen, const ulonglong2 *a, int lane)
{
	const int32_t k = r->as, cnt = r->cnt;
	const ulonglong2 a0 = a[k], a1 = a[k + cnt - 1];
	const int32_t q_span = (int32_t)(a0.y >> 32 & 0xff);
	const uint32_t rev = (uint32_t)(a0.x >> 63);
	int32_t blen = 0, mlen = 0;
	if (cnt > 0) {
		for (int i = k + 1 + lane; i < k + cnt; i += 32) {
			const ulonglong2 c = a[i], p = a[i - 1];
			const int span = (int)(c.y >> 32 & 0xff);
			const int tl = (int32_t)c.x - (int32_t)p.x;
			const int ql = (int32_t)c.y - (int32_t)p.y;
			blen += tl > ql ? tl : ql;
			mlen += tl > span && ql > span ? span : tl < ql ? tl : ql;
		}
		for (int d = 16; d; d >>= 1) { blen += __shfl_xor_sync(FULL, blen, d); mlen += __shfl_xor_sync(FULL, mlen, d); }
		blen += q_span; mlen += q_span;
	}
	if (lane == 0) {
		uint32_t bits = r->bits;
		r->bits = (bits & ~MM2B200_REG_REV) | (rev ? MM2B200_REG_REV : 0);
		r->rid = (int32_t)(a0.x << 1 >> 33);
		r->rs = (int32_t)a0.x + 1 > q_span ? (int32_t)a0.x + 1 - q_span : 0;
		r->re = (int32_t)a1.x + 1;
		if (!rev) {
			r->qs = (int32_t)a0.y + 1 - q_span;
			r->qe = (int32_t)a1.y + 1;
		} else {
			r->qs = qlen - ((int32_t)a1.y + 1);
			r->qe = qlen - ((int32_t)a0.y + 1 - q_span);
		}
		r->mlen = mlen; r->blen = blen;
	}
	__syncwarp();
}

// a correct ascending sort of n keys by one lane (the reference calls radix_sort_64 where equal keys are indistinguishable)
__device__ void sort_keys(uint64_t *k, uint64_t *junk, int n, int32_t *stk, int *s_head, int *s_tail, int lane)
{
	if (lane == 0 && n > 1) radix_sort_kv_exact(k, junk, n, stk, s_head, s_tail);
	__syncwarp();
}

// mm_set_sam_pri, hit.c:195-205
__device__ void set_sam_pri_u(int n, volatile Reg *r, int lane)
{
	int n_pri = 0;
	for (int i = 0; i < n; ++i) {
		const uint32_t bits = r[i].bits & ~MM2B200_REG_SAM_PRI;
		uint32_t nb = bits;
		if (r[i].id == r[i].parent) {
			++n_pri;
			if (n_pri == 1) nb |= MM2B200_REG_SAM_PRI;
		}
		if (lane == 0) r[i].bits = nb;
	}
	__syncwarp();
}

// mm_sync_regs, hit.c:207-228.  A parent id beyond the largest surviving id reads past tmp[] in the reference; here it is
// treated like a parent that was filtered out (PARENT_UNSET).
__device__ void sync_regs_u(int n_regs, volatile Reg *regs, volatile int32_t *tmp, int lane)
{
	if (n_regs <= 0) return;
	int max_id = -1;
	for (int i = 0; i < n_regs; ++i) max_id = max_id > regs[i].id ? max_id : regs[i].id;
	const int n_tmp = max_id + 1;
	for (int i = lane; i < n_tmp; i += 32) tmp[i] = -1;
	__syncwarp();
	if (lane == 0) for (int i = 0; i < n_regs; ++i) if (regs[i].id >= 0) tmp[regs[i].id] = i;
	__syncwarp();
	for (int i = lane; i < n_regs; i += 32) {
		const int parent = regs[i].parent;
		int np;
		if (parent == PARENT_TMP_PRI) np = i;
		else if (parent >= 0 && parent < n_tmp && tmp[parent] >= 0) np = tmp[parent];
		else np = PARENT_UNSET;
		regs[i].id = i; regs[i].parent = np;
	}
	__syncwarp();
	set_sam_pri_u(n_regs, regs, lane);
}

// mm_set_parent, hit.c:109-165 (mm_reg1_t.p is NULL on this path: the dp_max branch never runs)
__device__ void set_parent_u(float mask_level, int n, volatile Reg *r, volatile int32_t *w, uint64_t *cov, uint64_t *junk, int32_t *stk,
                             int *s_head, int *s_tail, int lane)
{
	if (n <= 0) return;
	for (int i = lane; i < n; i += 32) r[i].id = i;
	if (lane == 0) { w[0] = 0; r[0].parent = 0; }
	__syncwarp();
	int k = 1;
	for (int i = 1; i < n; ++i) {
		const int si = r[i].qs, ei = r[i].qe;
		int n_cov = 0, uncov_len = 0, j;
		for (j = 0; j < k; ++j) {
			const int wj = w[j];
			int sj = r[wj].qs, ej = r[wj].qe;
			if (ej <= si || sj >= ei) continue;
			if (sj < si) sj = si;
			if (ej > ei) ej = ei;
			if (lane == 0) cov[n_cov] = (uint64_t)(int64_t)sj << 32 | (uint64_t)(int64_t)ej;
			++n_cov;
		}
		j = k;
		if (n_cov > 0) {
			__syncwarp();
			sort_keys(cov, junk, n_cov, stk, s_head, s_tail, lane);
			int x = si;
			for (j = 0; j < n_cov; ++j) {
				const uint64_t c = ((volatile uint64_t*)cov)[j];
				if (c >> 32 > (uint64_t)(int64_t)x) uncov_len = (int)((uint64_t)(int64_t)uncov_len + ((c >> 32) - (uint64_t)(int64_t)x));
				x = (int32_t)c > x ? (int32_t)c : x;
			}
			if (ei > x) uncov_len += ei - x;
			for (j = 0; j < k; ++j) {
				const int wj = w[j];
				const int sj = r[wj].qs, ej = r[wj].qe;
				if (ej <= si || sj >= ei) continue;
				const int mn = ej - sj < ei - si ? ej - sj : ei - si;
				const int mx = ej - sj > ei - si ? ej - sj : ei - si;
				const int ol = si < sj ? (ei < sj ? 0 : ei < ej ? ei - sj : ej - sj) : (ej < si ? 0 : ej < ei ? ej - si : ei - si);
				const float lhs = __fsub_rn(__fdiv_rn(__int2float_rn(ol), __int2float_rn(mn)), __fdiv_rn(__int2float_rn(uncov_len), __int2float_rn(mx)));
				if (lhs > mask_level) {
					if (lane == 0) {
						const int sc = r[i].score, sub = r[wj].subsc;
						r[i].parent = r[wj].parent;
						r[wj].subsc = sub > sc ? sub : sc;
						if (r[i].cnt >= r[wj].cnt) r[wj].n_sub = r[wj].n_sub + 1;
					}
					break;
				}
			}
		}
		if (j == k) {
			if (lane == 0) { w[k] = i; r[i].parent = i; r[i].n_sub = 0; }
			++k;
		}
		__syncwarp();
	}
}

// mm_select_sub, hit.c:230-247: compacts r[] in place; r[p] is read from the partly compacted array exactly like the reference
__device__ int select_sub_u(float pri_ratio, int min_diff, int best_n, int n, Reg *r, volatile int32_t *tmp, int lane)
{
	if (!(pri_ratio > 0.0f) || n <= 0) return n;
	volatile Reg *vr = r;
	int k = 0, n_2nd = 0;
	for (int i = 0; i < n; ++i) {
		const int p = vr[i].parent;
		bool keep = false;
		if (p == i || (vr[i].bits & MM2B200_REG_INV)) keep = true;
		else {
			const int si = vr[i].score, sp = vr[p].score;
			if ((__int2float_rn(si) >= __fmul_rn(__int2float_rn(sp), pri_ratio) || si + min_diff >= sp) && n_2nd < best_n) {
				if (!(vr[i].qs == vr[p].qs && vr[i].qe == vr[p].qe && vr[i].rid == vr[p].rid && vr[i].rs == vr[p].rs && vr[i].re == vr[p].re))
					keep = true, ++n_2nd;
			}
		}
		if (keep) {
			if (k != i && lane == 0) copy_reg(r + k, r + i);
			++k;
			__syncwarp();
		}
	}
	if (k != n) sync_regs_u(k, vr, tmp, lane);
	return k;
}

// mm_filter_regs, hit.c:249-263 (p == NULL: only the min_cnt test applies)
__device__ int filter_regs_u(int min_cnt, int n, Reg *regs, int lane)
{
	volatile Reg *vr = regs;
	int k = 0;
	for (int i = 0; i < n; ++i) {
		const uint32_t bits = vr[i].bits;
		const bool flt = !(bits & MM2B200_REG_INV) && !(bits & MM2B200_REG_SEG_SPLIT) && vr[i].cnt < min_cnt;
		if (!flt) {
			if (k < i && lane == 0) copy_reg(regs + k, regs + i);
			++k;
			__syncwarp();
		}
	}
	return k;
}

// mm_squeeze_a, hit.c:265-283: the anchors of the surviving regions move to the front of a[], in the order of their offsets
__device__ void squeeze_a_u(int n_regs, volatile Reg *regs, ulonglong2 *a, uint64_t *aux, uint64_t *junk, int32_t *stk, int *s_head, int *s_tail, int lane)
{
	for (int i = lane; i < n_regs; i += 32) aux[i] = (uint64_t)(uint32_t)regs[i].as << 32 | (uint64_t)i;
	__syncwarp();
	sort_keys(aux, junk, n_regs, stk, s_head, s_tail, lane);
	int as = 0;
	for (int i = 0; i < n_regs; ++i) {
		const int idx = (int32_t)((volatile uint64_t*)aux)[i];
		const int src = regs[idx].as, cnt = regs[idx].cnt;
		if (src != as) {
			for (int off = 0; off < cnt; off += 32) {        // forward move, dst < src: a chunk is read whole before it is written
				ulonglong2 v;
				const bool in = off + lane < cnt;
				if (in) v = a[src + off + lane];
				__syncwarp();
				if (in) a[as + off + lane] = v;
				__syncwarp();
			}
			if (lane == 0) regs[idx].as = as;
		}
		as += cnt;
	}
	__syncwarp();
}

// mm_join_long, hit.c:290-345
__device__ int join_long_u(const mm2b200_post_params_t &opt, int qlen, int n_regs, Reg *R, ulonglong2 *a, uint64_t *aux, uint64_t *junk,
                           volatile int32_t *tmp, int32_t *stk, int *s_head, int *s_tail, int lane)
{
	volatile Reg *regs = R;
	if (n_regs < 2) return n_regs;
	squeeze_a_u(n_regs, regs, a, aux, junk, stk, s_head, s_tail, lane);
	int n_aux = 0, n_drop = 0;
	for (int i = 0; i < n_regs; ++i) {
		const int parent = regs[i].parent;
		if (parent == i || parent < 0) {
			if (lane == 0) aux[n_aux] = (uint64_t)(uint32_t)regs[i].as << 32 | (uint64_t)i;
			++n_aux;
		}
	}
	__syncwarp();
	sort_keys(aux, junk, n_aux, stk, s_head, s_tail, lane);
	for (int i = n_aux - 1; i >= 1; --i) {
		const int i0 = (int32_t)((volatile uint64_t*)aux)[i - 1], i1 = (int32_t)((volatile uint64_t*)aux)[i];
		volatile Reg *r0 = regs + i0, *r1 = regs + i1;
		const int as0 = r0->as, cnt0 = r0->cnt, as1 = r1->as, cnt1 = r1->cnt;
		if (as0 + cnt0 != as1) continue;
		if (r0->rid != r1->rid || (r0->bits & MM2B200_REG_REV) != (r1->bits & MM2B200_REG_REV)) continue;
		const ulonglong2 a0e = ld_anchor(a + as0 + cnt0 - 1), a1s = ld_anchor(a + as1);
		if (a1s.x <= a0e.x || (int32_t)a1s.y <= (int32_t)a0e.y) continue;
		int max_gap, min_gap;
		max_gap = min_gap = (int32_t)a1s.y - (int32_t)a0e.y;
		const uint64_t dx = a1s.x - a0e.x;
		max_gap = (int)((uint64_t)(int64_t)max_gap > dx ? (uint64_t)(int64_t)max_gap : dx);
		min_gap = (int)((uint64_t)(int64_t)min_gap < dx ? (uint64_t)(int64_t)min_gap : dx);
		if (max_gap > opt.max_join_long || min_gap > opt.max_join_short) continue;
		const float f = __fmul_rn(__fdiv_rn(__int2float_rn(opt.min_join_flank_sc), __int2float_rn(opt.max_join_long)), __int2float_rn(max_gap));
		const int sc_thres = __double2int_rz(__dadd_rn((double)f, .499));
		if (r0->score < sc_thres || r1->score < sc_thres) continue;
		if (r0->re - r0->rs < max_gap >> 1 || r0->qe - r0->qs < max_gap >> 1) continue;
		if (r1->re - r1->rs < max_gap >> 1 || r1->qe - r1->qs < max_gap >> 1) continue;
		if (lane == 0) {
			a[as1].y = a1s.y | SEED_LONG_JOIN;
			r0->cnt = cnt0 + cnt1; r0->score = r0->score + r1->score;
		}
		__syncwarp();
		set_coor_warp(r0, qlen, a, lane);
		if (lane == 0) { r1->cnt = 0; r1->parent = r0->id; }
		__syncwarp();
		++n_drop;
	}
	if (n_drop > 0) {
		for (int i = 0; i < n_regs; ++i) {
			const int parent = regs[i].parent;
			if (parent >= 0 && regs[i].id != parent) {
				const int pp = regs[parent].parent;
				if (pp >= 0 && pp != parent) { if (lane == 0) regs[i].parent = pp; }
			}
			__syncwarp();
		}
		n_regs = filter_regs_u(opt.min_cnt, n_regs, R, lane);
		sync_regs_u(n_regs, regs, tmp, lane);
	}
	return n_regs;
}

#define POST_WARPS 4
__global__ void __launch_bounds__(POST_WARPS * 32) k_post(PostWork w)
{
	__shared__ int s_ht[POST_WARPS][512];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	int *s_head = s_ht[wid], *s_tail = s_ht[wid] + 256;
	for (int r = blockIdx.x * POST_WARPS + wid; r < w.n_reads; r += gridDim.x * POST_WARPS) {
		const int n_u = w.n_u[r];
		if (n_u <= 0) { if (lane == 0) w.n_regs[r] = 0; continue; }
		const int64_t base = w.u_off[r];
		const uint64_t *u = w.u + base;
		ulonglong2 *a = w.b + w.b_off[r];
		Reg *R = w.R + base;
		uint64_t *zx = w.zx + base, *zy = w.zy + base, *cov = w.cov + base, *aux = w.aux + base;
		int32_t *wv = w.w + base, *tmp = w.tmp + base, *stk = w.stk + base;
		const uint32_t hash = w.hash[r];
		const int qlen = w.qlen[r];
		// ---- mm_gen_regs, hit.c:52-88
		int carry = 0;
		for (int i0 = 0; i0 < n_u; i0 += 32) {
			const int i = i0 + lane;
			const uint64_t ui = i < n_u ? u[i] : 0;
			const int cnt = (int32_t)ui;
			int incl = cnt;
			for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += t; }
			const int k = carry + incl - cnt;
			if (i < n_u) {
				const ulonglong2 f = a[k];
				const uint32_t h = (uint32_t)hash64((hash64(f.x) + hash64(f.y)) ^ hash);
				zx[i] = ui ^ h;
				zy[i] = (uint64_t)k << 32 | (uint64_t)(int64_t)cnt;
			}
			carry += __shfl_sync(FULL, incl, 31);
		}
		__syncwarp();
		if (lane == 0 && n_u > 1) radix_sort_kv_exact(zx, zy, n_u, stk, s_head, s_tail);     // radix_sort_128x: the order of equal keys matters
		__syncwarp();
		for (int i = lane; i < n_u; i += 32) {
			const uint64_t x = zx[n_u - 1 - i], y = zy[n_u - 1 - i];
			Reg g;
			memset(&g, 0, sizeof(g));
			g.id = i; g.parent = PARENT_UNSET;
			g.score = g.score0 = (int32_t)(x >> 32);
			g.hash = (uint32_t)x;
			g.cnt = (int32_t)y; g.as = (int32_t)(y >> 32);
			g.div = -1.0f;
			copy_reg(R + i, &g);
		}
		__syncwarp();
		for (int i = 0; i < n_u; ++i) set_coor_warp(R + i, qlen, a, lane);
		// ---- chain_post, map.c:238-247
		int n_regs = n_u;
		if (!(w.prm.flag & F_ALL_CHAINS)) {
			set_parent_u(w.prm.mask_level, n_regs, R, wv, cov, zx, stk, s_head, s_tail, lane);
			n_regs = select_sub_u(w.prm.pri_ratio, w.prm.min_diff, w.prm.best_n, n_regs, R, tmp, lane);
			if (!(w.prm.flag & (F_SPLICE | F_SR | F_NO_LJOIN)))
				n_regs = join_long_u(w.prm, qlen, n_regs, R, a, aux, zx, tmp, stk, s_head, s_tail, lane);
		}
		if (lane == 0) w.n_regs[r] = n_regs;
	}
}

// exclusive scan of n_regs over the reads by one CTA (the reads of a batch are few next to its anchors)
__global__ void __launch_bounds__(1024) k_post_scan(PostWork w)
{
	__shared__ long long s_warp[32];
	__shared__ long long s_carry;
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	if (tid == 0) s_carry = 0;
	__syncthreads();
	for (int r0 = 0; r0 < w.n_reads; r0 += 1024) {
		const int r = r0 + tid;
		const long long v = r < w.n_reads ? w.n_regs[r] : 0;
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
		if (r < w.n_reads) w.reg_off[r] = before;
		__syncthreads();
		if (tid == 1023) s_carry = before + v;
		__syncthreads();
	}
	if (tid == 0) { w.reg_off[w.n_reads] = s_carry; w.totals[0] = s_carry; }
}

// regions of read r: from the chain-indexed working array to their packed place
__global__ void __launch_bounds__(256) k_post_gather(PostWork w)
{
	const int lane = threadIdx.x & 31;
	const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
	for (int r = warp; r < w.n_reads; r += n_warps) {
		const int n = w.n_regs[r];
		const uint4 *src = (const uint4*)(w.R + w.u_off[r]);
		uint4 *dst = (uint4*)(w.out + w.reg_off[r]);
		for (int i = lane; i < n * 5; i += 32) dst[i] = src[i];
	}
}

inline size_t up256(size_t x) { return (x + 255) / 256 * 256; }

struct PostState {
	int device = 0;
	void *d_block = nullptr; size_t cap_chains = 0, cap_reads = 0;
	PostWork w{};
	// host API
	cudaStream_t stream = nullptr, copy_stream = nullptr;
	cudaEvent_t ev_ready[2] = { nullptr, nullptr }, ev_free[2] = { nullptr, nullptr };
	void *d_in = nullptr; size_t cap_in = 0;
	void *h_regs = nullptr, *h_off = nullptr, *h_a = nullptr, *h_aoff = nullptr, *h_tot = nullptr;
	size_t cap_h_regs = 0, cap_h_off = 0, cap_h_a = 0;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	double ms = 0; int64_t launches = 0;
	std::mutex mu;
};

void post_state_free(void *p)
{
	PostState *s = (PostState*)p;
	cudaSetDevice(s->device);
	if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
	if (s->copy_stream) { cudaStreamSynchronize(s->copy_stream); cudaStreamDestroy(s->copy_stream); }
	for (int k = 0; k < 2; ++k) { if (s->ev_ready[k]) cudaEventDestroy(s->ev_ready[k]); if (s->ev_free[k]) cudaEventDestroy(s->ev_free[k]); }
	if (s->d_block) cudaFree(s->d_block);
	if (s->d_in) cudaFree(s->d_in);
	for (void *h : { s->h_regs, s->h_off, s->h_a, s->h_aoff, s->h_tot }) if (h) cudaFreeHost(h);
	if (s->ev0) cudaEventDestroy(s->ev0);
	if (s->ev1) cudaEventDestroy(s->ev1);
	delete s;
}

std::mutex g_state_mu;
PostState *post_state(mm2b200_ctx_t *ctx)
{
	std::lock_guard<std::mutex> lk(g_state_mu);
	PostState *s = (PostState*)ctx_ext(ctx, 0);
	if (!s) {
		s = new PostState();
		s->device = ctx_device(ctx);
		ctx_set_ext(ctx, 0, s, post_state_free);
	}
	return s;
}

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

int ensure_work(PostState *s, size_t chains, size_t reads)
{
	if (chains <= s->cap_chains && reads <= s->cap_reads && s->d_block) return MM2B200_OK;
	size_t Cn = chains > s->cap_chains ? chains : s->cap_chains, Rn = reads > s->cap_reads ? reads : s->cap_reads;
	Cn += Cn / 8 + 1024; Rn += 64;
	if (s->d_block) { cudaDeviceSynchronize(); cudaFree(s->d_block); s->d_block = nullptr; }
	PostWork &w = s->w;
	struct Item { void **pp; size_t bytes; };
	Item items[] = {
		{ (void**)&w.zx, 8 * Cn }, { (void**)&w.zy, 8 * Cn }, { (void**)&w.cov, 8 * Cn }, { (void**)&w.aux, 8 * Cn },
		{ (void**)&w.w, 4 * Cn }, { (void**)&w.tmp, 4 * Cn }, { (void**)&w.stk, 4 * Cn + 256 },
		{ (void**)&w.R, sizeof(Reg) * Cn }, { (void**)&w.out, sizeof(Reg) * Cn },
		{ (void**)&w.n_regs, 4 * Rn }, { (void**)&w.reg_off, 8 * (Rn + 1) }, { (void**)&w.totals, 64 },
	};
	size_t total = 0;
	for (auto &it : items) total += up256(it.bytes);
	if (cudaMalloc(&s->d_block, total) != cudaSuccess) {
		cudaGetLastError();
		s->cap_chains = s->cap_reads = 0;
		return failf(MM2B200_ERR_NOMEM, "cudaMalloc(%zu bytes) for the post-processing of %zu chains failed", total, Cn);
	}
	char *cur = (char*)s->d_block;
	for (auto &it : items) { *it.pp = cur; cur += up256(it.bytes); }
	s->cap_chains = Cn; s->cap_reads = Rn;
	return MM2B200_OK;
}

int check_post(const mm2b200_post_params_t *pp)
{
	if (!pp) return failf(MM2B200_ERR_ARG, "post-processing parameters are NULL");
	if (pp->best_n < 0 || pp->max_join_long <= 0 || pp->min_cnt < 0) return failf(MM2B200_ERR_ARG, "bad post-processing parameters (best_n %d, max_join_long %d, min_cnt %d)", pp->best_n, pp->max_join_long, pp->min_cnt);
	return MM2B200_OK;
}

int pin_reserve(void **p, size_t *cap, size_t bytes)
{
	if (*p && bytes <= *cap) return MM2B200_OK;
	if (*p) { cudaFreeHost(*p); *p = nullptr; *cap = 0; }
	const size_t want = bytes + bytes / 4 + 4096;
	if (cudaHostAlloc(p, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu bytes) failed in the post-processing stage", want); }
	*cap = want;
	return MM2B200_OK;
}

} // namespace

// n_chains_cap: an upper bound of the batch's chains (the caller knows it: every chain has >= 1 anchor)
static int post_device_impl(mm2b200_ctx_t *ctx, PostState *s, const mm2b200_post_params_t *post, int64_t n_reads, const mm2b200_dev_result_t *chains,
                            int64_t n_chains_cap, const uint32_t *d_hash, const int32_t *d_qlen, cudaStream_t st, mm2b200_dev_regs_t *out)
{
	int rc;
	if ((rc = ensure_work(s, (size_t)n_chains_cap + 1, (size_t)n_reads + 1))) return rc;
	PostWork &w = s->w;
	w.n_u = chains->d_n_u; w.u_off = chains->d_u_off; w.b_off = chains->d_b_off; w.u = chains->d_u; w.b = (ulonglong2*)chains->d_b;
	w.hash = d_hash; w.qlen = d_qlen; w.n_reads = (int32_t)n_reads; w.prm = *post;
	const bool prof = ctx_profile(ctx) != 0;
	if (prof) {
		if (!s->ev0) { CU_TRY(cudaEventCreate(&s->ev0)); CU_TRY(cudaEventCreate(&s->ev1)); }
		CU_TRY(cudaEventRecord(s->ev0, st));
	}
	int launches = 0;
	if (n_reads > 0) {
		const int sm = ctx_sm_count(ctx);
		int grid = (int)((n_reads + POST_WARPS - 1) / POST_WARPS);
		if (grid > sm * 16) grid = sm * 16;
		k_post<<<grid, POST_WARPS * 32, 0, st>>>(w);
		k_post_scan<<<1, 1024, 0, st>>>(w);
		int g2 = (int)((n_reads + 7) / 8);
		if (g2 > sm * 8) g2 = sm * 8;
		k_post_gather<<<g2, 256, 0, st>>>(w);
		launches = 3;
	} else {
		CU_TRY(cudaMemsetAsync(w.reg_off, 0, 8, st));
		CU_TRY(cudaMemsetAsync(w.totals, 0, 64, st));
	}
	CU_TRY(cudaGetLastError());
	if (prof) CU_TRY(cudaEventRecord(s->ev1, st));
	s->launches += launches;
	ctx_add_launches(ctx, launches);
	out->d_reg_off = w.reg_off; out->d_regs = w.out; out->d_n_regs = w.n_regs; out->d_totals = w.totals;
	return MM2B200_OK;
}

extern "C" int mm2b200_post_device(mm2b200_ctx_t *ctx, const mm2b200_post_params_t *post, int64_t n_reads, const mm2b200_dev_result_t *chains,
                                   int64_t n_chains_cap, const uint32_t *d_hash, const int32_t *d_qlen, void *stream, mm2b200_dev_regs_t *out)
{
	int rc;
	if (!ctx || !chains || !out) return failf(MM2B200_ERR_ARG, "NULL ctx, chains or result");
	if ((rc = check_post(post))) return rc;
	if (n_reads < 0 || n_chains_cap < 0 || (n_reads && (!d_hash || !d_qlen))) return failf(MM2B200_ERR_ARG, "bad post-processing arguments");
	PostState *s = post_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	CU_TRY(cudaSetDevice(s->device));
	return post_device_impl(ctx, s, post, n_reads, chains, n_chains_cap, d_hash, d_qlen, (cudaStream_t)stream, out);
}

// after the caller has synchronised the stream: add the events' span to the stage time (the composed path, seed_api.cu)
namespace mm2b200 {
void post_stage_account(mm2b200_ctx *ctx)
{
	PostState *s = post_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	if (ctx_profile(ctx) && s->ev0) {
		float t = 0;
		if (cudaEventElapsedTime(&t, s->ev0, s->ev1) == cudaSuccess) s->ms += t;
		cudaGetLastError();
	}
}
}

extern "C" int mm2b200_post_stage_ms(mm2b200_ctx_t *ctx, double *ms, int64_t *launches)
{
	if (!ctx) return failf(MM2B200_ERR_ARG, "NULL context");
	PostState *s = post_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	if (ms) *ms = s->ms;
	if (launches) *launches = s->launches;
	return MM2B200_OK;
}

extern "C" int mm2b200_chain_regs_batch(mm2b200_ctx_t *ctx, const mm2b200_params_t *prm, const mm2b200_post_params_t *post,
                                        int64_t n_reads, const int64_t *offs, const mm128_t *a, const uint32_t *hash, const int32_t *qlen,
                                        int want_anchors, mm2b200_regs_t *res)
{
	int rc;
	if (!ctx || !res || !prm) return failf(MM2B200_ERR_ARG, "NULL ctx, params or result");
	if ((rc = check_post(post))) return rc;
	if (prm->n_segs > 1) return failf(MM2B200_ERR_ARG, "post-processing of multi-segment reads (n_segs = %d) is outside this path", prm->n_segs);
	if (n_reads < 0 || (n_reads && (!offs || !hash || !qlen))) return failf(MM2B200_ERR_ARG, "bad batch arguments");
	for (int64_t r = 0; r < n_reads; ++r) if (offs[r + 1] < offs[r]) return failf(MM2B200_ERR_ARG, "offs is not non-decreasing at read %lld", (long long)r);
	const int64_t a0 = n_reads ? offs[0] : 0, A = n_reads ? offs[n_reads] - a0 : 0;
	if (A && !a) return failf(MM2B200_ERR_ARG, "NULL anchor array");
	PostState *s = post_state(ctx);
	std::lock_guard<std::mutex> lk(s->mu);
	memset(res, 0, sizeof(*res));
	CU_TRY(cudaSetDevice(s->device));
	if (!s->stream) {
		CU_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
		CU_TRY(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
		for (int k = 0; k < 2; ++k) {
			CU_TRY(cudaEventCreateWithFlags(&s->ev_ready[k], cudaEventDisableTiming));
			CU_TRY(cudaEventCreateWithFlags(&s->ev_free[k], cudaEventDisableTiming));
		}
	}
	cudaStream_t st = s->stream, cs = s->copy_stream;
	const size_t R = (size_t)n_reads;
	// Chunks of whole reads of about 8 Mi anchors: the anchors of chunk c+1 cross PCIe (copy stream, second input buffer) while
	// chunk c is chained and post-processed; only the regions (80 B each) come back, so the bus carries little else.
	int64_t chunk_anchors = (int64_t)8 << 20;
	{ const char *e = getenv("MM2B200_REGS_CHUNK"); const long long v = e ? atoll(e) : 0; if (v > 0) chunk_anchors = (int64_t)v; }
	std::vector<int64_t> cut(1, 0);
	int64_t max_a = 1;
	for (int64_t r = 0; r < n_reads; ++r)
		if (offs[r + 1] - offs[cut.back()] >= chunk_anchors && r + 1 < n_reads) cut.push_back(r + 1);
	cut.push_back(n_reads);
	const size_t n_chunks = cut.size() - 1;
	for (size_t c = 0; c < n_chunks; ++c) { const int64_t na = offs[cut[c + 1]] - offs[cut[c]]; if (na > max_a) max_a = na; }
	if (max_a >= ((int64_t)1 << 31) - 64) return failf(MM2B200_ERR_ARG, "a read of %lld anchors exceeds the 2^31 index space of one device batch", (long long)max_a);
	// device inputs: two anchor buffers, per-chunk offsets (rebased), hash, qlen
	const size_t buf_b = up256(16 * ((size_t)max_a + 1));
	const size_t need = 2 * buf_b + up256(8 * (R + n_chunks + 1)) + 2 * up256(4 * (R + 1));
	if (need > s->cap_in) {
		if (s->d_in) { CU_TRY(cudaDeviceSynchronize()); cudaFree(s->d_in); s->d_in = nullptr; s->cap_in = 0; }
		const size_t want = need + need / 4;
		if (cudaMalloc(&s->d_in, want) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaMalloc(%zu bytes) for chunks of %lld anchors failed", want, (long long)max_a); }
		s->cap_in = want;
	}
	char *cur = (char*)s->d_in;
	mm128_t *d_a[2] = { (mm128_t*)cur, (mm128_t*)(cur + buf_b) }; cur += 2 * buf_b;
	int64_t *d_offs = (int64_t*)cur; cur += up256(8 * (R + n_chunks + 1));
	uint32_t *d_hash = (uint32_t*)cur; cur += up256(4 * (R + 1));
	int32_t *d_qlen = (int32_t*)cur;
	if ((rc = pin_reserve(&s->h_off, &s->cap_h_off, 8 * (R + n_chunks + 1) * 2 + 16 * (R + 1) + 64))) return rc;
	int64_t *h_off = (int64_t*)s->h_off;                         // chunk c's offsets, relative to its first anchor, at h_off + cut[c] + c
	int64_t *h_co = h_off + (R + n_chunks + 1);                  // per chunk: reg_off as the device wrote it
	int64_t *h_roff = h_co + (R + n_chunks + 1);                 // reg_off of the whole batch
	for (size_t c = 0; c < n_chunks; ++c)
		for (int64_t r = cut[c]; r <= cut[c + 1]; ++r) h_off[r + (int64_t)c] = offs[r] - offs[cut[c]];
	// the per-read arrays travel first on the copy stream, from page-locked memory: a copy from pageable memory queued on the
	// compute stream can end up behind the anchor uploads on the copy engine and hold the first chunk's kernels back
	uint32_t *h_hash = (uint32_t*)(h_roff + (R + 1));
	int32_t *h_qlen = (int32_t*)(h_hash + (R + 1));
	if (R) { memcpy(h_hash, hash, 4 * R); memcpy(h_qlen, qlen, 4 * R); }
	CU_TRY(cudaMemcpyAsync(d_offs, h_off, 8 * (R + n_chunks + 1), cudaMemcpyHostToDevice, cs));
	if (R) {
		CU_TRY(cudaMemcpyAsync(d_hash, h_hash, 4 * R, cudaMemcpyHostToDevice, cs));
		CU_TRY(cudaMemcpyAsync(d_qlen, h_qlen, 4 * R, cudaMemcpyHostToDevice, cs));
	}
	if (!s->h_tot) CU_TRY(cudaHostAlloc(&s->h_tot, 128, cudaHostAllocDefault));
	int64_t *tot = (int64_t*)s->h_tot;
	if ((rc = pin_reserve(&s->h_regs, &s->cap_h_regs, sizeof(Reg) * (2 * R + 64)))) return rc;
	int64_t *h_aoff = nullptr;
	if (want_anchors) {
		if ((rc = pin_reserve(&s->h_a, &s->cap_h_a, up256(16 * ((size_t)A + 1)) + 8 * (R + 1) + 256))) return rc;
		h_aoff = (int64_t*)((char*)s->h_a + up256(16 * ((size_t)A + 1)));
		res->a = (const mm128_t*)s->h_a; res->a_off = h_aoff;
	}
	auto upload = [&](size_t c) -> int {                          // chunk c's anchors into buffer c % 2, once its previous user is done
		const int b = (int)(c & 1);
		const int64_t lo = offs[cut[c]], na = offs[cut[c + 1]] - lo;
		if (c >= 2) CU_TRY(cudaStreamWaitEvent(cs, s->ev_free[b], 0));
		if (na) CU_TRY(cudaMemcpyAsync(d_a[b], a + lo, 16 * (size_t)na, cudaMemcpyHostToDevice, cs));
		CU_TRY(cudaEventRecord(s->ev_ready[b], cs));
		return MM2B200_OK;
	};
	int64_t tot_regs = 0, tot_chains = 0, tot_b = 0;
	if ((rc = upload(0))) return rc;
	for (size_t c = 0; c < n_chunks; ++c) {
		const int b = (int)(c & 1);
		const int64_t r0 = cut[c], nr = cut[c + 1] - cut[c], na = offs[cut[c + 1]] - offs[cut[c]];
		CU_TRY(cudaStreamWaitEvent(st, s->ev_ready[b], 0));
		mm2b200_dev_result_t dev;
		if ((rc = mm2b200_chain_device(ctx, prm, nr, na, d_offs + r0 + (int64_t)c, d_a[b], st, &dev))) return rc;
		const int64_t cap = na / (prm->min_cnt > 1 ? prm->min_cnt : 1) + 1;   // every chain holds at least max(min_cnt, 1) anchors
		mm2b200_dev_regs_t dr;
		if ((rc = post_device_impl(ctx, s, post, nr, &dev, cap, d_hash + r0, d_qlen + r0, st, &dr))) return rc;
		CU_TRY(cudaEventRecord(s->ev_free[b], st));
		CU_TRY(cudaMemcpyAsync(tot, dr.d_totals, 64, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaMemcpyAsync(tot + 8, dev.d_totals, 64, cudaMemcpyDeviceToHost, st));
		CU_TRY(cudaMemcpyAsync(h_co + r0 + (int64_t)c, dr.d_reg_off, 8 * ((size_t)nr + 1), cudaMemcpyDeviceToHost, st));
		if (want_anchors) CU_TRY(cudaMemcpyAsync(h_aoff + r0, dev.d_b_off, 8 * ((size_t)nr + 1), cudaMemcpyDeviceToHost, st));
		if (c + 1 < n_chunks && (rc = upload(c + 1))) return rc;  // the next chunk's upload is queued before this chunk is waited for
		CU_TRY(cudaStreamSynchronize(st));
		const int64_t n_regs = tot[0], n_b = tot[8 + 1];
		if (sizeof(Reg) * (size_t)(tot_regs + n_regs + 1) > s->cap_h_regs) {        // grow, keeping what earlier chunks delivered
			const size_t want = sizeof(Reg) * (size_t)(tot_regs + n_regs + 1) * 2;
			void *q = nullptr;
			if (cudaHostAlloc(&q, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return failf(MM2B200_ERR_NOMEM, "cudaHostAlloc(%zu bytes) for the regions failed", want); }
			memcpy(q, s->h_regs, sizeof(Reg) * (size_t)tot_regs);
			cudaFreeHost(s->h_regs);
			s->h_regs = q; s->cap_h_regs = want;
		}
		if (n_regs) CU_TRY(cudaMemcpyAsync((Reg*)s->h_regs + tot_regs, dr.d_regs, sizeof(Reg) * (size_t)n_regs, cudaMemcpyDeviceToHost, st));
		if (want_anchors && n_b) CU_TRY(cudaMemcpyAsync((mm128_t*)s->h_a + tot_b, dev.d_b, 16 * (size_t)n_b, cudaMemcpyDeviceToHost, st));
		for (int64_t r = 0; r < nr; ++r) h_roff[r0 + r] = tot_regs + h_co[r0 + (int64_t)c + r];
		if (want_anchors) for (int64_t r = 0; r < nr; ++r) h_aoff[r0 + r] += tot_b;
		if (ctx_profile(ctx) && s->ev0) {
			float t = 0;
			if (cudaEventElapsedTime(&t, s->ev0, s->ev1) == cudaSuccess) s->ms += t;
			cudaGetLastError();
		}
		tot_regs += n_regs; tot_chains += tot[8 + 0]; tot_b += n_b;
	}
	h_roff[R] = tot_regs;
	if (want_anchors) h_aoff[R] = tot_b;
	CU_TRY(cudaStreamSynchronize(st));
	CU_TRY(cudaStreamSynchronize(cs));
	tot[8 + 0] = tot_chains;
	res->n_reads = n_reads; res->reg_off = h_roff; res->regs = (const mm2b200_reg_t*)s->h_regs; res->n_chains = tot[8 + 0];
	return MM2B200_OK;
}
