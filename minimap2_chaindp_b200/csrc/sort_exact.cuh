// sort_exact.cuh -- the reference's (unstable) radix sort replayed on the device; shared by the chain tail
// (chain_kernels.cu) and the seeding kernels (seed_kernels.cu).
#pragma once
#include <stdint.h>

#define RS_MIN 64   // RS_MIN_SIZE, ksort.h:98

// rs_insertsort (ksort.h:105-115) on parallel arrays (key, payload)
static __device__ void ins_sort_kv(uint64_t *kx, uint64_t *ky, int beg, int end)
{
	for (int i = beg + 1; i < end; ++i) {
		if (kx[i] < kx[i - 1]) {
			const uint64_t tx = kx[i], ty = ky[i];
			int j = i;
			for (; j > beg && tx < kx[j - 1]; --j) { kx[j] = kx[j - 1]; ky[j] = ky[j - 1]; }
			kx[j] = tx; ky[j] = ty;
		}
	}
}

// radix_sort_128x (ksort.h:116-151, misc.c:136-137) replayed step for step by one thread so that the
// (unstable) order of equal keys matches the reference.  Sub-buckets are independent, so the recursion is
// replaced by an explicit stack of (beg, end, shift) in stk[].
static __device__ void radix_sort_kv_exact(uint64_t *kx, uint64_t *ky, int n, int32_t *stk, int *head, int *tail)
{
	if (n <= RS_MIN) { ins_sort_kv(kx, ky, 0, n); return; }
	int sp = 0;
	stk[0] = 0; stk[1] = n; stk[2] = 56; sp = 3;
	while (sp) {
		sp -= 3;
		const int beg = stk[sp], end = stk[sp + 1];
		int shift = stk[sp + 2];
		// A level at which every key of the range has the same digit leaves the range as it is and hands all of it to the next
		// level (one bucket, larger than RS_MIN like the range itself): such levels are skipped -- straight to the highest digit
		// in which two keys differ (none: the keys are all equal and every remaining level is the identity).  One pass of
		// independent loads instead of a counting pass and a permutation pass per skipped level.
		{
			const uint64_t x0 = kx[beg];
			uint64_t d = 0;
			for (int i = beg + 1; i < end; ++i) d |= kx[i] ^ x0;
			if (shift < 56) d &= ((uint64_t)1 << (shift + 8)) - 1;       // digits above this level are equal by construction
			if (d == 0) continue;
			shift = (63 - __clzll((long long)d)) & ~7;
		}
		for (int k = 0; k < 256; ++k) head[k] = 0;
		for (int i = beg; i < end; ++i) ++head[(kx[i] >> shift) & 0xff];
		for (int k = 0, acc = beg; k < 256; ++k) { const int c = head[k]; head[k] = acc; acc += c; tail[k] = acc; }
		for (int k = 0; k < 256;) {
			if (head[k] != tail[k]) {
				int l = (int)((kx[head[k]] >> shift) & 0xff);
				if (l != k) {
					uint64_t tx = kx[head[k]], ty = ky[head[k]];
					do {
						const uint64_t sx = tx, sy = ty;
						tx = kx[head[l]]; ty = ky[head[l]];
						kx[head[l]] = sx; ky[head[l]] = sy; ++head[l];
						l = (int)((tx >> shift) & 0xff);
					} while (l != k);
					kx[head[k]] = tx; ky[head[k]] = ty; ++head[k];
				} else ++head[k];
			} else ++k;
		}
		if (shift) {
			const int next = shift > 8 ? shift - 8 : 0;
			for (int k = 0, st = beg; k < 256; ++k) {
				const int en = tail[k], sz = en - st;
				if (sz > RS_MIN) { stk[sp] = st; stk[sp + 1] = en; stk[sp + 2] = next; sp += 3; }
				else if (sz > 1) ins_sort_kv(kx, ky, st, en);
				st = en;
			}
		}
	}
}
