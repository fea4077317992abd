/*
 * sketch_oracle.c -- CPU restatement of the reference's minimizer sketching.
 * TEST INFRASTRUCTURE ONLY: nothing in the product imports, links or executes this file.
 *
 * Restates mm_sketch (sketch.c:77-143) with its helpers (seq_nt4_table sketch.c:8-25, hash64 sketch.c:27-37, the 32-entry
 * queue of homopolymer run lengths sketch.c:39-57).  Parity: PINNED -- tests/test_sketch_oracle.py compares the output with
 * the compiled reference (oracle/_ref/libmm2ref_post.so: the reference's own sketch.c) on random, repetitive, ambiguous,
 * homopolymer-rich and palindromic sequences over the (w, k, HPC) settings of every preset.
 */
#include <stdint.h>
#include <string.h>
#include "chain_oracle.h"

static int nt4(unsigned char ch)                          /* seq_nt4_table, sketch.c:8-25 */
{
	switch (ch) {
	case 'A': case 'a': return 0;
	case 'C': case 'c': return 1;
	case 'G': case 'g': return 2;
	case 'T': case 't': case 'U': case 'u': return 3;
	default: return ch < 4 ? ch : 4;
	}
}

static uint64_t hash64m(uint64_t key, uint64_t mask)      /* sketch.c:27-37 */
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

#define PUSH(v) do { out[n_out++] = (v); } while (0)

/* out must hold len + 1 entries; returns the number of minimizers (appended from out[0]) */
int64_t orc_sketch(const char *str, int len, int w, int k, uint32_t rid, int is_hpc, orc_anchor_t *out)
{
	const uint64_t shift1 = 2 * (k - 1), mask = (1ULL << 2 * k) - 1, NONE = UINT64_MAX;
	uint64_t kmer[2] = { 0, 0 };
	int i, j, l = 0, buf_pos = 0, min_pos = 0, kmer_span = 0;
	int q[32], q_front = 0, q_count = 0;                  /* run lengths of the last k homopolymer-compressed bases */
	orc_anchor_t buf[256], min = { NONE, NONE };
	int64_t n_out = 0;
	memset(buf, 0xff, (size_t)w * 16);
	for (i = 0; i < len; ++i) {
		int c = nt4((unsigned char)str[i]);
		orc_anchor_t info = { NONE, NONE };
		if (c < 4) {
			int z;
			if (is_hpc) {
				int skip_len = 1;
				if (i + 1 < len && nt4((unsigned char)str[i + 1]) == c) {
					for (skip_len = 2; i + skip_len < len; ++skip_len)
						if (nt4((unsigned char)str[i + skip_len]) != c) break;
					i += skip_len - 1;                    /* i = last base of the run */
				}
				q[(q_count++ + q_front) & 0x1f] = skip_len;
				kmer_span += skip_len;
				if (q_count > k) { kmer_span -= q[q_front++]; q_front &= 0x1f; --q_count; }
			} else kmer_span = l + 1 < k ? l + 1 : k;
			kmer[0] = (kmer[0] << 2 | (uint64_t)c) & mask;
			kmer[1] = (kmer[1] >> 2) | (3ULL ^ (uint64_t)c) << shift1;
			if (kmer[0] == kmer[1]) continue;             /* strand unknown: the base leaves no trace in l, buf or buf_pos */
			z = kmer[0] < kmer[1] ? 0 : 1;
			++l;
			if (l >= k && kmer_span < 256) {
				info.x = hash64m(kmer[z], mask) << 8 | (uint64_t)kmer_span;
				info.y = (uint64_t)rid << 32 | (uint64_t)((uint32_t)i << 1) | (uint64_t)z;
			}
		} else l = 0, q_count = q_front = 0, kmer_span = 0;
		buf[buf_pos] = info;
		if (l == w + k - 1 && min.x != NONE) {            /* first full window: equal k-mers seen before it */
			for (j = buf_pos + 1; j < w; ++j) if (min.x == buf[j].x && buf[j].y != min.y) PUSH(buf[j]);
			for (j = 0; j < buf_pos; ++j) if (min.x == buf[j].x && buf[j].y != min.y) PUSH(buf[j]);
		}
		if (info.x <= min.x) {
			if (l >= w + k && min.x != NONE) PUSH(min);
			min = info, min_pos = buf_pos;
		} else if (buf_pos == min_pos) {
			if (l >= w + k - 1 && min.x != NONE) PUSH(min);
			for (j = buf_pos + 1, min.x = NONE; j < w; ++j) if (min.x >= buf[j].x) min = buf[j], min_pos = j;
			for (j = 0; j <= buf_pos; ++j) if (min.x >= buf[j].x) min = buf[j], min_pos = j;
			if (l >= w + k - 1 && min.x != NONE) {
				for (j = buf_pos + 1; j < w; ++j) if (min.x == buf[j].x && min.y != buf[j].y) PUSH(buf[j]);
				for (j = 0; j <= buf_pos; ++j) if (min.x == buf[j].x && min.y != buf[j].y) PUSH(buf[j]);
			}
		}
		if (++buf_pos == w) buf_pos = 0;
	}
	if (min.x != NONE) PUSH(min);
	return n_out;
}
