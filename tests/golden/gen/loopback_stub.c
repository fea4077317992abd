/*
 * loopback_stub.c -- stands in for the closed libfpgadrv.a (fpga.h:37-62) when generating golden vectors.
 *
 * The fork's software model of its device is fpga_work() (map.c:484): collect_seed_hits + mm_chain_dp_fpga on a
 * request packet.  This stub simply runs fpga_work synchronously on submit and queues the response packet for
 * fpga_get_retbuf, so the unmodified reference binary runs end to end on the CPU.
 * Test infrastructure only; used by tests/golden/make_golden.sh, never by the product.
 */
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

void *fpga_work(void *buf, int size, int *result_size);   /* map.c:484 */

typedef struct node { void *buf; int len; struct node *next; } node_t;
static node_t *head, *tail;
static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t cv = PTHREAD_COND_INITIALIZER;
static int exiting;

/* ---- optional dump of everything the device would receive (tests/golden/make_seed_golden.sh): $MM2_SEED_DUMP gets
 *   'P' int32[6]  bw, is_cdna, max_skip, min_sc, flag, max_occ            (fpga_set_params, main.c:243)
 *   'I' int32 type, int32 size, bytes                                      (fpga_load_index, main.c:201-204: B/H/V/P streams)
 *   'K' int32 size, bytes                                                  (one request packet, map.c:286-324)
 *   'A' int64 n, mm128_t[n]                                                (anchors of the packet's next read, map.c:523)
 *   'R' int32 size, bytes                                                  (the response packet fpga_work built, map.c:484-575)
 * With the dump on, packets are worked off one at a time so that the 'A' records follow their 'K'. */
#include <stdint.h>
#include <stdio.h>
typedef struct { uint64_t x, y; } stub_mm128_t;
extern void (*golden_on_fpga_input)(int64_t n, const stub_mm128_t *a) __attribute__((weak));   /* dump_hook.c; absent in the drop-in build */
static FILE *seed_fp;
static pthread_mutex_t seed_mu = PTHREAD_MUTEX_INITIALIZER, work_mu = PTHREAD_MUTEX_INITIALIZER;
static void seed_rec(char tag, const void *h, size_t hn, const void *b, size_t bn)
{
	if (!seed_fp) return;
	pthread_mutex_lock(&seed_mu);
	fputc(tag, seed_fp);
	if (hn) fwrite(h, 1, hn, seed_fp);
	if (bn) fwrite(b, 1, bn, seed_fp);
	fflush(seed_fp);
	pthread_mutex_unlock(&seed_mu);
}
static void on_fpga_input(int64_t n, const stub_mm128_t *a) { seed_rec('A', &n, 8, a, (size_t)n * 16); }
static void seed_open(void)
{
	const char *path = getenv("MM2_SEED_DUMP");
	if (path && !seed_fp && &golden_on_fpga_input && (seed_fp = fopen(path, "ab")) != 0) golden_on_fpga_input = on_fpga_input;
}

int fpga_init(int flag) { (void)flag; seed_open(); return 0; }
int fpga_init_noreset(int noblock) { (void)noblock; return 0; }
void fpga_finalize(void) {}
void fpga_set_block(void) {}
void fpga_set_params(int bw, int is_cdna, int max_skip, int min_sc, int flag, int max_occ)
{
	int32_t v[6] = { bw, is_cdna, max_skip, min_sc, flag, max_occ };
	seed_open();
	seed_rec('P', v, sizeof(v), 0, 0);
}
void fpga_load_index(void *addr, int size, int type)   /* fpga.h:62 */
{
	int32_t h[2] = { type, size };
	seed_open();
	seed_rec('I', h, sizeof(h), addr, (size_t)size);
}

void *fpga_get_writebuf(unsigned long size, int type) { (void)type; return malloc(size); }
void *fpga_get_writebuf_thread(unsigned long size, int type, int tid) { (void)type; (void)tid; return malloc(size); }

int fpga_writebuf_submit(void *addr, unsigned int size, unsigned int type)
{
	int len = 0;
	node_t *n = (node_t*)calloc(1, sizeof(node_t));
	(void)type;
	if (seed_fp) {
		int32_t sz = (int32_t)size;
		pthread_mutex_lock(&work_mu);
		seed_rec('K', &sz, 4, addr, size);
		n->buf = fpga_work(addr, (int)size, &len);
		sz = (int32_t)len;
		seed_rec('R', &sz, 4, n->buf, (size_t)len);
		pthread_mutex_unlock(&work_mu);
	} else n->buf = fpga_work(addr, (int)size, &len);
	n->len = len;
	free(addr);
	pthread_mutex_lock(&mu);
	if (tail) tail->next = n; else head = n;
	tail = n;
	pthread_cond_signal(&cv);
	pthread_mutex_unlock(&mu);
	return 0;
}

void *fpga_get_retbuf(int *len, int type)
{
	node_t *n;
	void *buf;
	(void)type;
	pthread_mutex_lock(&mu);
	while (!head && !exiting) pthread_cond_wait(&cv, &mu);
	if (!head) { pthread_mutex_unlock(&mu); *len = 0; return 0; }
	n = head; head = n->next; if (!head) tail = 0;
	pthread_mutex_unlock(&mu);
	buf = n->buf; *len = n->len; free(n);
	return buf;
}

int fpga_release_retbuf(void *addr) { free(addr); return 0; }

void fpga_exit_block(void)
{
	pthread_mutex_lock(&mu);
	exiting = 1;
	pthread_cond_broadcast(&cv);
	pthread_mutex_unlock(&mu);
}
