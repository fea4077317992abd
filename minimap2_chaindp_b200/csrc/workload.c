/*
 * workload.c -- deterministic synthetic anchor generator "G-anc" (SURVEY.md section 8d).
 *
 * Produces, for each synthetic read, the sorted mm128_t anchor array that the reference's collect_seed_hits
 * (map.c:187-236) would hand to mm_chain_dp: x = strand<<63 | rid<<32 | rpos, y = q_span<<32 | qpos
 * (map.c:217-229), sorted ascending by x (map.c:233).  One independent PRNG stream per read:
 * seed = splitmix64(base_seed) ^ read_id, so any subset of reads can be regenerated on any rank.
 *
 * Host-only C (no CUDA); built into libmm2workload.so by __graft_entry__.build().
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { uint64_t x, y; } wl_anchor_t;

typedef struct {
	uint64_t seed;          /* base seed (config id folded in by the caller) */
	int32_t read_len;       /* L: query length in bases */
	int32_t gap_max;        /* w: minimizer gaps ~ U{1..w} */
	double  p_survive;      /* probability a minimizer still matches: (1-e)^k */
	double  p_drift;        /* probability the diagonal drifts by +-1..3 at a minimizer (indels) */
	int32_t span_lo, span_hi; /* q_span ~ U{span_lo..span_hi} */
	int32_t n_contigs;      /* R reference sequences */
	int32_t contig_len;     /* length of each */
	double  p_background;   /* per minimizer: probability of a burst of random hits */
	double  bg_mean;        /* mean burst size (geometric) */
	double  p_repeat;       /* per minimizer: probability of a near-diagonal local repeat hit */
	int32_t repeat_radius;  /* +- bp around the true locus */
	int32_t repeat_band;    /* +- bp off the diagonal */
	int32_t n_loci;         /* >1: ava mode -- the read overlaps this many (target,strand) loci on average */
	int32_t reserved;
} wl_spec_t;

static inline uint64_t splitmix64(uint64_t *s)
{
	uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
	return z ^ (z >> 31);
}
static inline double u01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint32_t below(uint64_t *s, uint32_t n) { return n ? (uint32_t)(((splitmix64(s) >> 32) * (uint64_t)n) >> 32) : 0; }

static int cmp_x(const void *pa, const void *pb)
{
	const wl_anchor_t *a = (const wl_anchor_t*)pa, *b = (const wl_anchor_t*)pb;
	if (a->x != b->x) return a->x < b->x ? -1 : 1;
	return a->y < b->y ? -1 : a->y > b->y;
}

/* ascending by (x, y) -- the same total order as cmp_x, so the output is identical to qsort's (equal elements are
 * indistinguishable); long reads (map-pb, asm5: 5e4..4e6 anchors) spend most of the generator's time here, so they get
 * an LSD radix sort on x (16-bit digits, constant digits skipped) followed by an insertion pass over equal-x groups. */
static void sort_anchors(wl_anchor_t *a, int64_t n)
{
	wl_anchor_t *src = a, *dst, *tmp;
	uint32_t *cnt;
	int pass;
	int64_t i;
	if (n < 4096) { if (n > 1) qsort(a, (size_t)n, sizeof(wl_anchor_t), cmp_x); return; }
	tmp = (wl_anchor_t*)malloc((size_t)n * sizeof(wl_anchor_t));
	cnt = (uint32_t*)malloc(65536 * sizeof(uint32_t));
	dst = tmp;
	for (pass = 0; pass < 4; ++pass) {
		const int sh = 16 * pass;
		uint32_t acc = 0, k;
		memset(cnt, 0, 65536 * sizeof(uint32_t));
		for (i = 0; i < n; ++i) ++cnt[(src[i].x >> sh) & 0xffff];
		if (cnt[(src[0].x >> sh) & 0xffff] == (uint32_t)n) continue;      /* digit is the same everywhere */
		for (k = 0; k < 65536; ++k) { const uint32_t c = cnt[k]; cnt[k] = acc; acc += c; }
		for (i = 0; i < n; ++i) dst[cnt[(src[i].x >> sh) & 0xffff]++] = src[i];
		{ wl_anchor_t *t = src; src = dst; dst = t; }
	}
	if (src != a) memcpy(a, src, (size_t)n * sizeof(wl_anchor_t));
	for (i = 1; i < n; ++i)                                              /* ties on x: order by y */
		if (a[i].x == a[i - 1].x && a[i].y < a[i - 1].y) {
			wl_anchor_t v = a[i];
			int64_t j = i;
			while (j > 0 && a[j - 1].x == v.x && a[j - 1].y > v.y) { a[j] = a[j - 1]; --j; }
			a[j] = v;
		}
	free(cnt); free(tmp);
}

static inline void emit(wl_anchor_t *out, int64_t *k, uint32_t strand, uint32_t rid, uint32_t rpos, int32_t span, int32_t qpos)
{
	if (out) {
		out[*k].x = (uint64_t)strand << 63 | (uint64_t)rid << 32 | rpos;
		out[*k].y = (uint64_t)(uint32_t)span << 32 | (uint32_t)qpos;
	}
	++*k;
}

/* one colinear locus: query interval [q0,q1) laid on (strand, rid) starting at reference position r0 */
static void gen_locus(const wl_spec_t *sp, uint64_t *rng, wl_anchor_t *out, int64_t *k,
                      uint32_t strand, uint32_t rid, int64_t r0, int32_t q0, int32_t q1)
{
	int32_t q = q0;
	int64_t diag = r0 - q0;
	while (1) {
		int32_t span = sp->span_lo + (int32_t)below(rng, (uint32_t)(sp->span_hi - sp->span_lo + 1));
		q += 1 + (int32_t)below(rng, (uint32_t)sp->gap_max);
		if (q >= q1) break;
		if (u01(rng) < sp->p_drift) {
			int32_t d = 1 + (int32_t)below(rng, 3);
			diag += (splitmix64(rng) & 1) ? d : -d;
		}
		if (u01(rng) < sp->p_survive) {
			int64_t r = diag + q;
			if (r >= 0 && r < sp->contig_len) emit(out, k, strand, rid, (uint32_t)r, span, q);
		}
		if (u01(rng) < sp->p_background) {           /* burst of unrelated hits (repeat k-mers elsewhere) */
			double p = 1.0 / (sp->bg_mean > 1.0 ? sp->bg_mean : 1.0);
			do {
				emit(out, k, (uint32_t)(splitmix64(rng) & 1), below(rng, (uint32_t)sp->n_contigs),
				     below(rng, (uint32_t)sp->contig_len), span, q);
			} while (u01(rng) >= p);
		}
		if (u01(rng) < sp->p_repeat) {               /* local repeat: lands inside windows, exercises the skip logic */
			int64_t dq = (int64_t)below(rng, 2u * (uint32_t)sp->repeat_radius + 1) - sp->repeat_radius;
			int64_t off = (int64_t)below(rng, 2u * (uint32_t)sp->repeat_band + 1) - sp->repeat_band;
			int64_t r = diag + q + dq + off;
			if (r >= 0 && r < sp->contig_len) emit(out, k, strand, rid, (uint32_t)r, span, q);
		}
	}
}

static int64_t gen_read(const wl_spec_t *sp, int64_t read_id, wl_anchor_t *out)
{
	uint64_t s0 = sp->seed, rng;
	int64_t k = 0;
	rng = splitmix64(&s0) ^ (uint64_t)read_id;
	splitmix64(&rng);
	if (sp->n_loci <= 1) {
		uint32_t strand = (uint32_t)(splitmix64(&rng) & 1), rid = below(&rng, (uint32_t)sp->n_contigs);
		int64_t room = (int64_t)sp->contig_len - sp->read_len;
		int64_t r0 = room > 0 ? (int64_t)below(&rng, (uint32_t)room) : 0;
		gen_locus(sp, &rng, out, &k, strand, rid, r0, 0, sp->read_len);
	} else {
		/* ava: Poisson-ish number of overlapping targets (uniform in [n/2, 3n/2]); each overlap covers a
		 * prefix or suffix of the query of random length, on a random target at a random offset */
		int32_t n = sp->n_loci / 2 + (int32_t)below(&rng, (uint32_t)sp->n_loci + 1), l;
		for (l = 0; l < n; ++l) {
			uint32_t strand = (uint32_t)(splitmix64(&rng) & 1), rid = below(&rng, (uint32_t)sp->n_contigs);
			int32_t ov = 1 + (int32_t)below(&rng, (uint32_t)sp->read_len);
			int head = (int)(splitmix64(&rng) & 1);
			int32_t q0 = head ? 0 : sp->read_len - ov, q1 = head ? ov : sp->read_len;
			int64_t r0 = head ? (int64_t)sp->contig_len - ov : 0;
			if (r0 < 0) r0 = 0;
			gen_locus(sp, &rng, out, &k, strand, rid, r0, q0, q1);
		}
	}
	if (out && k > 1) sort_anchors(out, k);
	return k;
}

/* pass 1: offs[0..n_reads] for reads [first, first+n_reads).  Returns total anchors. */
int64_t wl_count(const wl_spec_t *sp, int64_t first, int64_t n_reads, int64_t *offs)
{
	int64_t r, tot = 0;
#pragma omp parallel for schedule(dynamic, 64)
	for (r = 0; r < n_reads; ++r) offs[r + 1] = gen_read(sp, first + r, 0);
	offs[0] = 0;
	for (r = 0; r < n_reads; ++r) { tot += offs[r + 1]; offs[r + 1] = tot; }
	return tot;
}

/* pass 2: fill a[offs[r]..offs[r+1]) for every read (same streams as pass 1) */
void wl_fill(const wl_spec_t *sp, int64_t first, int64_t n_reads, const int64_t *offs, wl_anchor_t *a)
{
	int64_t r;
#pragma omp parallel for schedule(dynamic, 64)
	for (r = 0; r < n_reads; ++r) gen_read(sp, first + r, a + offs[r]);
}

/* ---- driver-boundary pump (bench.py, row f2): what the fork's worker / receiver threads do with the driver symbols --------
 * n_threads submitters each take every n_threads-th packet: fpga_get_writebuf_thread (sleeping 50 us while it returns NULL,
 * map.c:439-441), memcpy, fpga_writebuf_submit; one receiver loops over fpga_get_retbuf / fpga_release_retbuf
 * (fpga_chaindp.c:241-262) until every packet is answered.  The symbols come in as function pointers so that this
 * generator library does not link the product.  Returns the wall time in seconds; *resp_bytes = bytes received. */
#include <pthread.h>
#include <time.h>
#include <unistd.h>

typedef void *(*wl_get_wbuf_t)(unsigned long, int, int);
typedef int (*wl_submit_t)(void*, unsigned int, unsigned int);
typedef void *(*wl_get_rbuf_t)(int*, int);
typedef int (*wl_release_t)(void*);

typedef struct {
	int tid, n_threads;
	int64_t n_packets;
	const int64_t *pkt_off;
	const char *pkts;
	wl_get_wbuf_t get_wbuf; wl_submit_t submit; wl_get_rbuf_t get_rbuf; wl_release_t release;
	int64_t got_bytes;
} wl_pump_t;

static void *wl_pump_send(void *arg)
{
	wl_pump_t *p = (wl_pump_t*)arg;
	for (int64_t i = p->tid; i < p->n_packets; i += p->n_threads) {
		const int64_t n = p->pkt_off[i + 1] - p->pkt_off[i];
		void *buf;
		while ((buf = p->get_wbuf((unsigned long)n, 0, p->tid)) == 0) usleep(50);
		memcpy(buf, p->pkts + p->pkt_off[i], (size_t)n);
		p->submit(buf, (unsigned int)n, 1);
	}
	return 0;
}

static void *wl_pump_recv(void *arg)
{
	wl_pump_t *p = (wl_pump_t*)arg;
	for (int64_t i = 0; i < p->n_packets; ++i) {
		int len = 0;
		void *buf = p->get_rbuf(&len, 3);
		if (len == 0) break;
		p->got_bytes += len;
		p->release(buf);
	}
	return 0;
}

double wl_shim_pump(int n_threads, int64_t n_packets, const int64_t *pkt_off, const char *pkts,
                    wl_get_wbuf_t get_wbuf, wl_submit_t submit, wl_get_rbuf_t get_rbuf, wl_release_t release, int64_t *resp_bytes)
{
	pthread_t th[65];
	wl_pump_t ctx[65];
	struct timespec t0, t1;
	if (n_threads < 1) n_threads = 1;
	if (n_threads > 64) n_threads = 64;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (int t = 0; t <= n_threads; ++t) {
		wl_pump_t c = { t, n_threads, n_packets, pkt_off, pkts, get_wbuf, submit, get_rbuf, release, 0 };
		ctx[t] = c;
		pthread_create(&th[t], 0, t < n_threads ? wl_pump_send : wl_pump_recv, &ctx[t]);
	}
	for (int t = 0; t <= n_threads; ++t) pthread_join(th[t], 0);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	if (resp_bytes) *resp_bytes = ctx[n_threads].got_bytes;
	return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
