/*
 * dump_hook.c -- records every call the reference makes to its chaining DP while it maps the fixture FASTA files.
 *
 * make_golden.sh compiles the reference's chain.c with -Dmm_chain_dp_fpga=ref_real_fpga
 * -Dmm_chain_dp_bottom=ref_real_bottom, so the symbols map.c calls (map.c:525,837,941) resolve to the wrappers
 * below, which append one binary record per call to $MM2_GOLDEN_DUMP and forward to the real functions.
 *
 *   'F' record: params[8] (max_dist_x,max_dist_y,bw,max_skip,min_sc,is_cdna,n_segs,0), int64 n, anchors[n],
 *               uint32 new_i, new_seed[new_i]
 *   'B' record: params[3] (min_cnt,min_sc,n_segs), uint32 new_i, new_seed[new_i], int32 n_u, u[n_u], int64 n_b, b[n_b]
 *
 * Test infrastructure only.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t x, y; } mm128_t;
struct new_seed { mm128_t seed; int32_t p; int32_t f; };

struct new_seed *ref_real_fpga(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_sc, int is_cdna, int n_segs,
                               int64_t n, mm128_t *a, uint32_t *_new_i);
mm128_t *ref_real_bottom(int min_cnt, int min_sc, int n_segs, int *n_u_, uint64_t **_u, void *km,
                         struct new_seed *fpga_a, uint32_t new_i);

static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;

/* seeding fixtures (make_seed_golden.sh): loopback_stub.c sets this to record the anchors collect_seed_hits (map.c:187)
 * hands to the DP, in the order of the reads of the request packet it is working on */
void (*golden_on_fpga_input)(int64_t n, const mm128_t *a) = 0;

static FILE *dump_open(void)
{
	const char *path = getenv("MM2_GOLDEN_DUMP");
	return path ? fopen(path, "ab") : 0;
}

struct new_seed *mm_chain_dp_fpga(int max_dist_x, int max_dist_y, int bw, int max_skip, int min_sc, int is_cdna, int n_segs,
                                  int64_t n, mm128_t *a, uint32_t *_new_i)
{
	mm128_t *copy = (mm128_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(mm128_t));
	struct new_seed *rec;
	FILE *fp;
	if (n > 0) memcpy(copy, a, (size_t)n * sizeof(mm128_t));
	if (golden_on_fpga_input) golden_on_fpga_input(n, a);
	rec = ref_real_fpga(max_dist_x, max_dist_y, bw, max_skip, min_sc, is_cdna, n_segs, n, a, _new_i);   /* frees a */
	pthread_mutex_lock(&mu);
	if ((fp = dump_open()) != 0) {
		int32_t prm[8] = { max_dist_x, max_dist_y, bw, max_skip, min_sc, is_cdna, n_segs, 0 };
		fputc('F', fp);
		fwrite(prm, 4, 8, fp);
		fwrite(&n, 8, 1, fp);
		fwrite(copy, sizeof(mm128_t), (size_t)n, fp);
		fwrite(_new_i, 4, 1, fp);
		fwrite(rec, sizeof(struct new_seed), *_new_i, fp);
		fclose(fp);
	}
	pthread_mutex_unlock(&mu);
	free(copy);
	return rec;
}

mm128_t *mm_chain_dp_bottom(int min_cnt, int min_sc, int n_segs, int *n_u_, uint64_t **_u, void *km,
                            struct new_seed *fpga_a, uint32_t new_i)
{
	mm128_t *b = ref_real_bottom(min_cnt, min_sc, n_segs, n_u_, _u, km, fpga_a, new_i);
	FILE *fp;
	pthread_mutex_lock(&mu);
	if ((fp = dump_open()) != 0) {
		int32_t prm[3] = { min_cnt, min_sc, n_segs }, n_u = *n_u_;
		int64_t n_b = 0;
		int i;
		for (i = 0; i < n_u; ++i) n_b += (int32_t)(*_u)[i];
		fputc('B', fp);
		fwrite(prm, 4, 3, fp);
		fwrite(&new_i, 4, 1, fp);
		fwrite(fpga_a, sizeof(struct new_seed), new_i, fp);
		fwrite(&n_u, 4, 1, fp);
		if (n_u) fwrite(*_u, 8, (size_t)n_u, fp);
		fwrite(&n_b, 8, 1, fp);
		if (n_b) fwrite(b, sizeof(mm128_t), (size_t)n_b, fp);
		fclose(fp);
	}
	pthread_mutex_unlock(&mu);
	return b;
}
