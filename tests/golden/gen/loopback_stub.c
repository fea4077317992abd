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

int fpga_init(int flag) { (void)flag; return 0; }
int fpga_init_noreset(int noblock) { (void)noblock; return 0; }
void fpga_finalize(void) {}
void fpga_set_block(void) {}
void fpga_set_params(int bw, int is_cdna, int max_skip, int min_sc, int flag, int max_occ)
{ (void)bw; (void)is_cdna; (void)max_skip; (void)min_sc; (void)flag; (void)max_occ; }
int fpga_load_index(int type, const char *buf, int size) { (void)type; (void)buf; (void)size; return 0; }

void *fpga_get_writebuf(unsigned long size, int type) { (void)type; return malloc(size); }
void *fpga_get_writebuf_thread(unsigned long size, int type, int tid) { (void)type; (void)tid; return malloc(size); }

int fpga_writebuf_submit(void *addr, unsigned int size, unsigned int type)
{
	int len = 0;
	node_t *n = (node_t*)calloc(1, sizeof(node_t));
	(void)type;
	n->buf = fpga_work(addr, (int)size, &len);
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
