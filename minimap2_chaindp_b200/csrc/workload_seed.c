/*
 * workload_seed.c -- deterministic synthetic workload for the seeding stage, "G-idx" (DESIGN.md, row f1).
 *
 * There are no sequences here: a reference is a list of minimizer SITES (hash value, sequence, position, strand), which is
 * all the index ever sees of it.  The generator
 *   1. draws the sites (gaps ~ U{1..w}, hash uniform over 2k bits, a fraction copied from earlier sites so that some
 *      minimizers have several occurrences, and a few "hot" values with hundreds),
 *   2. lays the index out in the four streams the fork ships to its device (index.c:603-720: B bucket records, H hash
 *      tables in 64-byte groups of 8 slots, V values, P position lists; positions packed refid 21 | refpos 21 | strand 1 |
 *      rank 21) with khash's own probing (khash.h:218-231; hash = key>>1, index.c:23), and
 *   3. samples reads as windows of consecutive sites: a site survives sequencing errors with probability p_match (its
 *      hash is kept), otherwise it is replaced by a random value; query positions follow the reference with indel drift;
 *      half of the reads are reverse-complemented.
 * One PRNG stream per read (seed ^ read id), so any rank can generate any range of reads.
 *
 * Host-only C; built into libmm2workload.so.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
	uint64_t seed;
	int64_t n_sites;
	int32_t hash_bits;      /* 2k */
	int32_t b;              /* log2 buckets (mi->b, 14 in minimap2) */
	int32_t gap_max;        /* w */
	int32_t ref_len;        /* bases per reference sequence, < 2^21 (the fork packs refpos in 21 bits) */
	double  p_dup;          /* a site repeats the hash of a random earlier site */
	double  p_hot;          /* a site takes one of n_hot hot values */
	int32_t n_hot;
	int32_t reserved;
} wls_spec_t;

typedef struct {
	wls_spec_t sp;
	uint64_t *hash;         /* [n_sites] */
	uint64_t *pos;          /* [n_sites] packed position */
	uint8_t *B, *H, *V, *P;
	int64_t nB, nH, nV, nP; /* bytes */
	int64_t n_keys, max_occ_seen;
} wls_t;

static inline uint64_t sm64(uint64_t *s)
{
	uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
	return z ^ (z >> 31);
}
static inline double sm01(uint64_t *s) { return (double)(sm64(s) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint64_t smbelow(uint64_t *s, uint64_t n) { return n ? (uint64_t)(((__uint128_t)sm64(s) * n) >> 64) : 0; }

typedef struct { uint64_t h; int64_t i; } hs_t;
static int cmp_hs(const void *a, const void *b)
{
	const hs_t *x = (const hs_t*)a, *y = (const hs_t*)b;
	if (x->h != y->h) return x->h < y->h ? -1 : 1;
	return x->i < y->i ? -1 : x->i > y->i;
}

void wls_free(wls_t *w)
{
	if (!w) return;
	free(w->hash); free(w->pos); free(w->B); free(w->H); free(w->V); free(w->P); free(w);
}

static void wls_layout(wls_t *w);

wls_t *wls_build(const wls_spec_t *sp)
{
	const int64_t N = sp->n_sites;
	wls_t *w = (wls_t*)calloc(1, sizeof(wls_t));
	uint64_t rng = sp->seed ^ 0x1D5EEDULL, hmask = sp->hash_bits >= 64 ? ~0ULL : (1ULL << sp->hash_bits) - 1;
	int64_t i;
	if (!w) return 0;
	w->sp = *sp;
	w->hash = (uint64_t*)malloc(8 * (size_t)(N + 1));
	w->pos = (uint64_t*)malloc(8 * (size_t)(N + 1));
	uint64_t *hot = (uint64_t*)malloc(8 * (size_t)(sp->n_hot + 1));
	for (i = 0; i < sp->n_hot; ++i) hot[i] = sm64(&rng) & hmask;
	{	/* 1. sites */
		uint32_t refid = 0, refpos = 0;
		for (i = 0; i < N; ++i) {
			const double u = sm01(&rng);
			refpos += 1 + (uint32_t)smbelow(&rng, (uint64_t)sp->gap_max);
			if (refpos >= (uint32_t)sp->ref_len) { ++refid; refpos = 1 + (uint32_t)smbelow(&rng, (uint64_t)sp->gap_max); }
			if (u < sp->p_hot && sp->n_hot > 0) w->hash[i] = hot[smbelow(&rng, (uint64_t)sp->n_hot)];
			else if (u < sp->p_hot + sp->p_dup && i > 0) w->hash[i] = w->hash[smbelow(&rng, (uint64_t)i)];
			else w->hash[i] = sm64(&rng) & hmask;
			w->pos[i] = (uint64_t)(refid & 0x1fffff) << 43 | (uint64_t)(refpos & 0x1fffff) << 22 | (sm64(&rng) & 1) << 21 | (refid & 0x1fffff);
		}
	}
	free(hot);
	wls_layout(w);
	return w;
}

/* 2. the four streams from w->hash[] / w->pos[] */
static void wls_layout(wls_t *w)
{
	const wls_spec_t *sp = &w->sp;
	const int64_t N = sp->n_sites;
	int64_t i;
	/* group by (bucket, key): sort by (hash with the bucket bits on top) so that buckets come out in order */
	hs_t *hs = (hs_t*)malloc(sizeof(hs_t) * (size_t)(N + 1));
	const uint64_t bmask = (1ULL << sp->b) - 1;
	for (i = 0; i < N; ++i) { const uint64_t h = w->hash[i]; hs[i].h = (h & bmask) << (64 - sp->b) | h >> sp->b; hs[i].i = i; }
	qsort(hs, (size_t)N, sizeof(hs_t), cmp_hs);
	const int64_t n_bkt = 1LL << sp->b;
	w->nB = 16 * n_bkt;
	w->B = (uint8_t*)calloc(1, (size_t)w->nB);
	/* sizes first */
	int64_t tot_slots = 0, tot_p = 0, at = 0;
	int64_t *bk_beg = (int64_t*)malloc(8 * (size_t)(n_bkt + 1));
	for (int64_t k = 0; k < n_bkt; ++k) {
		bk_beg[k] = at;
		while (at < N && (int64_t)(hs[at].h >> (64 - sp->b)) == k) ++at;
	}
	bk_beg[n_bkt] = at;
	int64_t *slot0 = (int64_t*)malloc(8 * (size_t)n_bkt), *p0 = (int64_t*)malloc(8 * (size_t)n_bkt);
	uint32_t *nbk = (uint32_t*)malloc(4 * (size_t)n_bkt);
	for (int64_t k = 0; k < n_bkt; ++k) {
		int64_t keys = 0, multi = 0, j = bk_beg[k];
		while (j < bk_beg[k + 1]) {
			int64_t e = j + 1;
			while (e < bk_beg[k + 1] && hs[e].h == hs[j].h) ++e;
			++keys; if (e - j > 1) multi += e - j;
			if (e - j > w->max_occ_seen) w->max_occ_seen = e - j;
			j = e;
		}
		uint32_t nb = 0;
		if (keys) { nb = 4; while ((double)keys >= 0.77 * nb) nb <<= 1; }      /* khash's upper bound, khash.h __ac_HASH_UPPER */
		nbk[k] = nb; slot0[k] = tot_slots; p0[k] = tot_p;
		tot_slots += (nb + 7) / 8 * 8; tot_p += multi;
		w->n_keys += keys;
	}
	w->nH = tot_slots / 8 * 64; w->nV = tot_slots * 8; w->nP = tot_p * 8;
	w->H = (uint8_t*)calloc(1, (size_t)w->nH + 64);
	w->V = (uint8_t*)calloc(1, (size_t)w->nV + 64);
	w->P = (uint8_t*)calloc(1, (size_t)w->nP + 64);
	#pragma omp parallel for schedule(dynamic, 64)
	for (int64_t k = 0; k < n_bkt; ++k) {
		const uint32_t nb = nbk[k];
		uint64_t rec[2] = { 0, 0 };
		if (nb) {
			const int64_t s0 = slot0[k];
			uint64_t *V = (uint64_t*)w->V + s0, *P = (uint64_t*)w->P + p0[k];
			uint8_t *fl = (uint8_t*)malloc(nb);      /* 2 = empty, 0 = live */
			int64_t j = bk_beg[k], np = 0;
			memset(fl, 2, nb);
			rec[1] = (uint64_t)s0 << 28 | (uint64_t)p0[k] >> 8;                          /* index.c:622 */
			rec[0] = ((uint64_t)p0[k] & 0xff) << 56 | (uint64_t)nb << 24;                /* index.c:623 */
			while (j < bk_beg[k + 1]) {
				int64_t e = j + 1;
				while (e < bk_beg[k + 1] && hs[e].h == hs[j].h) ++e;
				const uint64_t hi = w->hash[hs[j].i] >> sp->b;                         /* key >> 1 */
				uint32_t s = (uint32_t)hi & (nb - 1), step = 0;
				while (fl[s] == 0) s = (s + (++step)) & (nb - 1);                       /* khash.h:226 */
				fl[s] = 0;
				uint64_t key = hi << 1, val;
				if (e - j == 1) { key |= 1; val = w->pos[hs[j].i]; }                    /* index.c: single occurrence lives in the value */
				else { val = (uint64_t)np << 32 | (uint64_t)(e - j); for (int64_t q = j; q < e; ++q) P[np++] = w->pos[hs[q].i]; }
				uint8_t *grp = w->H + 64 * ((s0 + s) >> 3);
				memcpy(grp + 4 + 6 * ((s0 + s) & 7), &key, 6);
				V[s] = val;
				j = e;
			}
			for (uint32_t g = 0; g < (nb + 7) / 8; ++g) {                               /* flags word of slots (8g & ~15) .., index.c:661-664 */
				uint32_t word = 0, base16 = (g * 8) & ~15u;
				for (uint32_t q = 0; q < 16; ++q) word |= (uint32_t)(base16 + q < nb ? fl[base16 + q] : 2) << (2 * q);
				memcpy(w->H + 64 * ((s0 >> 3) + g), &word, 4);
			}
			free(fl);
		}
		memcpy(w->B + 16 * k, rec, 16);
	}
	free(hs); free(bk_beg); free(slot0); free(p0); free(nbk);
}

/* An index over caller-supplied sites (hash = minimizer value, pos = packed position): what mm_idx_gen would build from the
 * sketch of a reference (index.c:340-416), laid out like above.  Takes copies. */
wls_t *wls_build_sites(int32_t b, int32_t hash_bits, int64_t n_sites, const uint64_t *hash, const uint64_t *pos)
{
	wls_t *w = (wls_t*)calloc(1, sizeof(wls_t));
	if (!w) return 0;
	w->sp.n_sites = n_sites; w->sp.b = b; w->sp.hash_bits = hash_bits;
	w->hash = (uint64_t*)malloc(8 * (size_t)(n_sites + 1));
	w->pos = (uint64_t*)malloc(8 * (size_t)(n_sites + 1));
	memcpy(w->hash, hash, 8 * (size_t)n_sites);
	memcpy(w->pos, pos, 8 * (size_t)n_sites);
	wls_layout(w);
	return w;
}

void wls_streams(const wls_t *w, const uint8_t **B, int64_t *nB, const uint8_t **H, int64_t *nH, const uint8_t **V, int64_t *nV,
                 const uint8_t **P, int64_t *nP, int64_t *n_keys, int64_t *max_occ)
{
	*B = w->B; *nB = w->nB; *H = w->H; *nH = w->nH; *V = w->V; *nV = w->nV; *P = w->P; *nP = w->nP;
	if (n_keys) *n_keys = w->n_keys;
	if (max_occ) *max_occ = w->max_occ_seen;
}

/* ---- reads ------------------------------------------------------------------------------------------------------ */
typedef struct {
	uint64_t seed;
	int32_t read_len;       /* bases */
	int32_t span;           /* k */
	double p_match;         /* a site's k-mer survives the read's errors: (1-e)^k */
	double p_drift;         /* indel drift of the query position per site */
} wls_reads_t;

/* one read: walks sites from a random start; returns the number of minimizers; writes them if mv != NULL
 * (x = hash<<8 | span, y = pos<<1 | strand, ascending query position) */
static int64_t wls_one(const wls_t *w, const wls_reads_t *rs, int64_t read_id, uint64_t *mv, int32_t *qlen, uint32_t *bid)
{
	uint64_t rng = rs->seed ^ (uint64_t)read_id * 0xD1B54A32D192ED03ULL;
	const int64_t N = w->sp.n_sites;
	const uint64_t hmask = w->sp.hash_bits >= 64 ? ~0ULL : (1ULL << w->sp.hash_bits) - 1;
	int64_t s = (int64_t)smbelow(&rng, (uint64_t)N), n = 0;
	const int rev = (int)(sm64(&rng) & 1);
	const uint64_t ref0 = w->pos[s] >> 43;
	const int64_t p0 = (int64_t)((w->pos[s] >> 22) & 0x1fffff);
	int64_t drift = 0;
	const int64_t L = rs->read_len;
	/* count first so that reversed reads can be written back to front */
	int64_t e = s;
	while (e < N && (w->pos[e] >> 43) == ref0 && (int64_t)((w->pos[e] >> 22) & 0x1fffff) - p0 + rs->span + 8 < L) ++e;
	n = e - s;
	if (qlen) *qlen = (int32_t)L;
	if (bid) *bid = (uint32_t)((ref0 + 1 + smbelow(&rng, 7)) & 0x1fffff);     /* a rank that is not the target's own */
	if (!mv) return n;
	int64_t last_q = -1;
	for (int64_t j = 0; j < n; ++j) {
		const int64_t site = s + j;
		uint64_t h = w->hash[site];
		if (sm01(&rng) >= rs->p_match) h = sm64(&rng) & hmask;
		if (sm01(&rng) < rs->p_drift) drift += (int64_t)smbelow(&rng, 7) - 3;
		int64_t q = (int64_t)((w->pos[site] >> 22) & 0x1fffff) - p0 + rs->span - 1 + 4 + drift;   /* position of the k-mer's last base */
		if (q <= last_q) q = last_q + 1;
		if (q >= L) q = L - 1;
		last_q = q;
		const uint64_t strand = ((w->pos[site] >> 21) & 1) ^ (uint64_t)rev;
		const int64_t qq = rev ? L - 1 - q + rs->span - 1 : q;
		const int64_t slot = rev ? n - 1 - j : j;
		mv[2 * slot] = h << 8 | (uint64_t)rs->span;
		mv[2 * slot + 1] = (uint64_t)qq << 1 | strand;
	}
	return n;
}

/* mv_off[n_reads+1] (relative to the first read), qlen[], bid[] */
int64_t wls_reads_count(const wls_t *w, const wls_reads_t *rs, int64_t first, int64_t n_reads, int64_t *mv_off, int32_t *qlen, uint32_t *bid)
{
	#pragma omp parallel for schedule(static, 256)
	for (int64_t r = 0; r < n_reads; ++r) mv_off[r + 1] = wls_one(w, rs, first + r, 0, qlen + r, bid + r);
	mv_off[0] = 0;
	for (int64_t r = 0; r < n_reads; ++r) mv_off[r + 1] += mv_off[r];
	return mv_off[n_reads];
}

void wls_reads_fill(const wls_t *w, const wls_reads_t *rs, int64_t first, int64_t n_reads, const int64_t *mv_off, uint64_t *mv)
{
	#pragma omp parallel for schedule(dynamic, 64)
	for (int64_t r = 0; r < n_reads; ++r) wls_one(w, rs, first + r, mv + 2 * mv_off[r], 0, 0);
}


/* ---- "G-seq": a synthetic genome and reads sampled from it, as BASES (rows f4 + f1 + DP + f3 composed) ------------------ */
typedef struct {
	uint64_t seed;
	int32_t n_contigs, contig_len;      /* contig_len < 2^21 (the fork packs refpos in 21 bits) */
	double  p_repeat;                   /* share of the genome overwritten by copies of earlier 2-6 kb segments */
	int32_t read_len;                   /* bases taken from the genome per read, before errors */
	double  err;                        /* per-base error rate, one third each substitution / insertion / deletion */
} wlg_spec_t;

/* genome[n_contigs * contig_len], contig c at c * contig_len */
void wlg_genome(const wlg_spec_t *sp, char *genome)
{
	const int64_t G = (int64_t)sp->n_contigs * sp->contig_len;
	#pragma omp parallel for schedule(static)
	for (int64_t c = 0; c < sp->n_contigs; ++c) {
		uint64_t rng = sp->seed ^ (0x6E0ULL + (uint64_t)c) * 0x9E3779B97F4A7C15ULL;
		char *g = genome + c * sp->contig_len;
		for (int32_t i = 0; i < sp->contig_len; i += 32) {
			uint64_t v = sm64(&rng);
			for (int32_t j = 0; j < 32 && i + j < sp->contig_len; ++j, v >>= 2) g[i + j] = "ACGT"[v & 3];
		}
	}
	{	/* repeats: serial, so that a copy can itself be copied */
		uint64_t rng = sp->seed ^ 0x4E9EA7ULL;
		int64_t todo = (int64_t)(sp->p_repeat * (double)G);
		while (todo > 0) {
			const int64_t len = 2000 + (int64_t)smbelow(&rng, 4000);
			const int64_t cs = (int64_t)smbelow(&rng, (uint64_t)sp->n_contigs), cd = (int64_t)smbelow(&rng, (uint64_t)sp->n_contigs);
			if (len + 1 >= sp->contig_len) break;
			const int64_t ps = (int64_t)smbelow(&rng, (uint64_t)(sp->contig_len - len)), pd = (int64_t)smbelow(&rng, (uint64_t)(sp->contig_len - len));
			memmove(genome + cd * sp->contig_len + pd, genome + cs * sp->contig_len + ps, (size_t)len);
			todo -= len;
		}
	}
}

/* one read; returns its length; writes the bases if out != NULL */
static int64_t wlg_one(const wlg_spec_t *sp, const char *genome, int64_t read_id, char *out, uint32_t *bid)
{
	uint64_t rng = sp->seed ^ (uint64_t)(read_id + 1) * 0xD1B54A32D192ED03ULL;
	const int64_t c = (int64_t)smbelow(&rng, (uint64_t)sp->n_contigs);
	const int64_t L = sp->read_len < sp->contig_len ? sp->read_len : sp->contig_len;
	const int64_t p = (int64_t)smbelow(&rng, (uint64_t)(sp->contig_len - L + 1));
	const int rev = (int)(sm64(&rng) & 1);
	const char *g = genome + c * sp->contig_len + p;
	const double e3 = sp->err / 3.0;
	int64_t n = 0;
	if (bid) *bid = (uint32_t)((c + 1 + (int64_t)smbelow(&rng, 7)) % (sp->n_contigs > 1 ? sp->n_contigs : 1));
	else (void)smbelow(&rng, 7);
	for (int64_t i = 0; i < L; ++i) {
		const int64_t src = rev ? L - 1 - i : i;
		char ch = g[src];
		if (rev) ch = ch == 'A' ? 'T' : ch == 'C' ? 'G' : ch == 'G' ? 'C' : 'A';
		const double u = sm01(&rng);
		if (u < e3) continue;                                                   /* deletion */
		if (u < 2 * e3) ch = "ACGT"[sm64(&rng) & 3];                             /* substitution (may be silent) */
		else if (u < 3 * e3) { const char ins = "ACGT"[sm64(&rng) & 3]; if (out) out[n] = ins; ++n; }   /* insertion in front (drawn in the counting pass too) */
		if (out) out[n] = ch;
		++n;
	}
	return n;
}

int64_t wlg_reads_count(const wlg_spec_t *sp, const char *genome, int64_t first, int64_t n_reads, int64_t *seq_off, uint32_t *bid)
{
	#pragma omp parallel for schedule(static, 64)
	for (int64_t r = 0; r < n_reads; ++r) seq_off[r + 1] = wlg_one(sp, genome, first + r, 0, bid ? bid + r : 0);
	seq_off[0] = 0;
	for (int64_t r = 0; r < n_reads; ++r) seq_off[r + 1] += seq_off[r];
	return seq_off[n_reads];
}

void wlg_reads_fill(const wlg_spec_t *sp, const char *genome, int64_t first, int64_t n_reads, const int64_t *seq_off, char *seq)
{
	#pragma omp parallel for schedule(dynamic, 64)
	for (int64_t r = 0; r < n_reads; ++r) wlg_one(sp, genome, first + r, seq + seq_off[r], 0);
}
