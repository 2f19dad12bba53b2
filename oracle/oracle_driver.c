/*
 * oracle/oracle_driver.c — TEST INFRASTRUCTURE, not product code.
 *
 * Batch/threads driver around the two LZ4E entry points
 * (LZ4E_compress_default, LZ4E_decompress_safe; reference prototypes at
 * lz4e/include/lz4e.h:47-51).  It is linked twice:
 *   - with the reference's own sources  -> oracle/_ref/liblz4e_ref.so
 *   - with oracle/lz4e_port.c           -> oracle/_build/liblz4e_port.so
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
 * load either library; the product never does.
 *
 * One block per task, one thread per requested core, each thread with its
 * own LZ4E_MEM_COMPRESS-byte wrkmem.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <linux/bio.h>

/* benchmark data generator (tools/synth), so that bench.py's CPU arm needs no
 * other native library than this one */
#include "../tools/synth/lz4e_synth.h"

#define ORACLE_WRKMEM_BYTES 17440 /* LZ4E_MEM_COMPRESS, lz4e/include/lz4e.h:17-22 */

int LZ4E_compress_default(const struct bio_vec *src, struct bio_vec *dst,
			  struct bvec_iter *srcIter, struct bvec_iter *dstIter,
			  void *wrkmem);
int LZ4E_decompress_safe(const char *source, char *dest, int compressedSize,
			 int maxDecompressedSize);

struct job {
	/* compress */
	const struct bio_vec *src_bvecs;
	const uint64_t *src_first;
	const struct bvec_iter *src_iters;
	struct bio_vec *dst_bvecs;
	const uint64_t *dst_first;
	const struct bvec_iter *dst_iters;
	struct bvec_iter *src_iters_out;
	struct bvec_iter *dst_iters_out;
	/* decompress */
	const char *dsrc_base;
	const uint64_t *dsrc_off;
	const int *dsrc_len;
	char *ddst_base;
	const uint64_t *ddst_off;
	const int *ddst_cap;
	/* common */
	int64_t nblocks;
	int *rets;
	int64_t next; /* atomic work counter */
	int decompress;
};

static void *worker(void *arg)
{
	struct job *j = (struct job *)arg;
	void *wrkmem = NULL;

	if (!j->decompress) {
		if (posix_memalign(&wrkmem, 64, ORACLE_WRKMEM_BYTES))
			return NULL;
		memset(wrkmem, 0xA5, ORACLE_WRKMEM_BYTES); /* callee must zero it */
	}
	for (;;) {
		int64_t b = __atomic_fetch_add(&j->next, 1, __ATOMIC_RELAXED);
		if (b >= j->nblocks)
			break;
		if (j->decompress) {
			j->rets[b] = LZ4E_decompress_safe(
				j->dsrc_base + j->dsrc_off[b],
				j->ddst_base + j->ddst_off[b], j->dsrc_len[b],
				j->ddst_cap[b]);
		} else {
			struct bvec_iter si = j->src_iters[b];
			struct bvec_iter di = j->dst_iters[b];

			j->rets[b] = LZ4E_compress_default(
				j->src_bvecs + j->src_first[b],
				j->dst_bvecs + j->dst_first[b], &si, &di, wrkmem);
			if (j->src_iters_out)
				j->src_iters_out[b] = si;
			if (j->dst_iters_out)
				j->dst_iters_out[b] = di;
		}
	}
	free(wrkmem);
	return NULL;
}

static double run_job(struct job *j, int nthreads)
{
	struct timespec t0, t1;
	pthread_t *th;
	int i;

	if (nthreads < 1)
		nthreads = 1;
	th = (pthread_t *)calloc((size_t)nthreads, sizeof(*th));
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (i = 1; i < nthreads; i++)
		pthread_create(&th[i], NULL, worker, j);
	worker(j);
	for (i = 1; i < nthreads; i++)
		pthread_join(th[i], NULL);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	free(th);
	return (double)(t1.tv_sec - t0.tv_sec) +
	       1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/*
 * Block b compresses src_bvecs[src_first[b]...] under iterator src_iters[b]
 * into dst_bvecs[dst_first[b]...] under dst_iters[b].  Iterators are copied,
 * the end state is returned through *_iters_out when non-NULL.
 * Returns wall seconds of the threaded region.
 */
double oracle_compress_batch(const struct bio_vec *src_bvecs,
			     const uint64_t *src_first,
			     const struct bvec_iter *src_iters,
			     struct bio_vec *dst_bvecs, const uint64_t *dst_first,
			     const struct bvec_iter *dst_iters, int64_t nblocks,
			     int *rets, struct bvec_iter *src_iters_out,
			     struct bvec_iter *dst_iters_out, int nthreads)
{
	struct job j;

	memset(&j, 0, sizeof(j));
	j.src_bvecs = src_bvecs;
	j.src_first = src_first;
	j.src_iters = src_iters;
	j.dst_bvecs = dst_bvecs;
	j.dst_first = dst_first;
	j.dst_iters = dst_iters;
	j.src_iters_out = src_iters_out;
	j.dst_iters_out = dst_iters_out;
	j.nblocks = nblocks;
	j.rets = rets;
	return run_job(&j, nthreads);
}

/* Contiguous decode of block b: src_base+src_off[b] (src_len[b] bytes) into
 * dst_base+dst_off[b] (capacity dst_cap[b]). */
double oracle_decompress_batch(const char *src_base, const uint64_t *src_off,
			       const int *src_len, char *dst_base,
			       const uint64_t *dst_off, const int *dst_cap,
			       int64_t nblocks, int *rets, int nthreads)
{
	struct job j;

	memset(&j, 0, sizeof(j));
	j.decompress = 1;
	j.dsrc_base = src_base;
	j.dsrc_off = src_off;
	j.dsrc_len = src_len;
	j.ddst_base = dst_base;
	j.ddst_off = dst_off;
	j.ddst_cap = dst_cap;
	j.nblocks = nblocks;
	j.rets = rets;
	return run_job(&j, nthreads);
}

struct synth_job {
	uint8_t *base;
	uint64_t stride, nblocks, first;
	uint32_t n;
	int64_t next;
};

static void *synth_worker(void *arg)
{
	struct synth_job *j = (struct synth_job *)arg;

	for (;;) {
		int64_t b = __atomic_fetch_add(&j->next, 64, __ATOMIC_RELAXED);
		int64_t e = b + 64;

		if (b >= (int64_t)j->nblocks)
			break;
		if (e > (int64_t)j->nblocks)
			e = (int64_t)j->nblocks;
		for (; b < e; b++)
			lz4e_synth_block(j->base + (uint64_t)b * j->stride, j->n,
					 (uint32_t)((j->first + (uint64_t)b) & 3u),
					 (uint32_t)(j->first + (uint64_t)b));
	}
	return NULL;
}

/* synthetic benchmark contents: nblocks blocks of n bytes at base + b * stride,
 * class (first + b) & 3, seed first + b */
void oracle_synth_blocks(uint8_t *base, uint64_t stride, uint32_t n, uint64_t nblocks,
			 uint64_t first, int nthreads)
{
	struct synth_job j = { base, stride, nblocks, first, n, 0 };
	pthread_t th[256];
	int i;

	if (nthreads < 1)
		nthreads = 1;
	if (nthreads > 256)
		nthreads = 256;
	for (i = 1; i < nthreads; i++)
		pthread_create(&th[i], NULL, synth_worker, &j);
	synth_worker(&j);
	for (i = 1; i < nthreads; i++)
		pthread_join(th[i], NULL);
}

/* layout probes so the Python side can assert its ctypes mirror is right */
int oracle_sizeof_bio_vec(void) { return (int)sizeof(struct bio_vec); }
int oracle_sizeof_bvec_iter(void) { return (int)sizeof(struct bvec_iter); }
int oracle_sizeof_page(void) { return (int)sizeof(struct page); }
