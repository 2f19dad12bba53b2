/*
 * lz4e_chunk.c — the lz4e_bdev chunk layer and request path in userspace, on
 * the GPU LZ4E library.  Interface and behaviour follow the reference's
 * lz4e_bdev/lz4e_chunk.c:19-159 and lz4e_bdev/lz4e_req.c:53-213 (see
 * include/lz4e_chunk.h); the implementation is batch-first.
 */
#include <errno.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/lz4e.h"
#include "../../include/lz4e_chunk.h"
#include "lz4e_internal.h"

/* ---- buffer pool: pinned pages with a device mirror ------------------------- */

#define POOL_CLASSES 513 /* free lists for runs of 1..512 pages */

/* free runs are linked OUT of band (page index -> next page index + 1): pool
 * pages are arena pages, and a batch brings the touched byte range of an arena
 * back from the device in one bulk copy, which may overwrite free pages that
 * lie inside that range */

/*
 * One pool per CONTEXT (= per GPU), found through a small registry; a thread
 * chooses the context its chunk calls run on with lz4e_chunk_bind(), so one
 * process can drive several GPUs through the chunk layer (one submitter thread
 * per GPU).  Threads that never bind use the context of the first
 * lz4e_chunk_setup(), else the default context.
 */
#define MAX_POOLS 16

struct pool {
	lz4e_b200_ctx *ctx;
	lz4e_b200_arena *arena;
	uint8_t *base;
	size_t npages, bump;
	size_t freel[POOL_CLASSES]; /* first free run of that many pages, + 1 */
	size_t *next;		    /* per page index */
	size_t live;
};

static pthread_mutex_t p_mu = PTHREAD_MUTEX_INITIALIZER; /* registry and every pool */
static struct pool *g_pools[MAX_POOLS];
static lz4e_b200_ctx *g_first_ctx;	   /* context of the first lz4e_chunk_setup() */
static __thread lz4e_b200_ctx *t_ctx;	   /* lz4e_chunk_bind() of this thread */

static void pool_release_locked(int i)
{
	struct pool *pl = g_pools[i];

	g_pools[i] = NULL;
	if (g_first_ctx == pl->ctx)
		g_first_ctx = NULL;
	lz4e_b200_arena_destroy(pl->arena);
	free(pl->next);
	free(pl);
}

/* the pool's context is being destroyed: the pool goes with it (chunks still
 * alive keep dangling buffers — a caller bug, reported) */
static void pool_ctx_dtor(lz4e_b200_ctx *ctx, void *p)
{
	int i;

	(void)p;
	pthread_mutex_lock(&p_mu);
	for (i = 0; i < MAX_POOLS; i++)
		if (g_pools[i] && g_pools[i]->ctx == ctx) {
			if (g_pools[i]->live)
				fprintf(stderr,
					"lz4e_chunk: context destroyed with %zu pool buffers live\n",
					g_pools[i]->live);
			pool_release_locked(i);
		}
	pthread_mutex_unlock(&p_mu);
}

/* the context this thread's chunk calls run on */
static lz4e_b200_ctx *current_ctx_locked(void)
{
	if (t_ctx)
		return t_ctx;
	if (g_first_ctx)
		return g_first_ctx;
	return lz4e_b200_default_ctx();
}

/* the pool of ctx, created on first use; NULL when that fails */
static struct pool *pool_get_locked(lz4e_b200_ctx *ctx, size_t pool_pages)
{
	struct pool *pl;
	int i, slot = -1;

	if (!ctx)
		return NULL;
	for (i = 0; i < MAX_POOLS; i++) {
		if (g_pools[i] && g_pools[i]->ctx == ctx)
			return g_pools[i];
		if (!g_pools[i] && slot < 0)
			slot = i;
	}
	if (slot < 0)
		return NULL;
	if (!pool_pages) {
		const char *e = getenv("LZ4E_CHUNK_POOL_MB");
		long mb = e ? atol(e) : 1024;

		if (mb < 1)
			mb = 1;
		pool_pages = (size_t)mb << (20 - PAGE_SHIFT);
	}
	pl = calloc(1, sizeof(*pl));
	if (!pl)
		return NULL;
	pl->next = calloc(pool_pages, sizeof(*pl->next));
	if (!pl->next || lz4e_b200_arena_create(ctx, pool_pages, &pl->arena) != LZ4E_B200_OK) {
		free(pl->next);
		free(pl);
		return NULL;
	}
	pl->ctx = ctx;
	pl->base = (uint8_t *)lz4e_b200_arena_pages(pl->arena);
	pl->npages = pool_pages;
	lz4e_ctx_attach(ctx, LZ4E_ATTACH_CHUNK_POOL, 8, pool_ctx_dtor);
	g_pools[slot] = pl;
	return pl;
}

/* the pool whose pages hold p, or NULL */
static struct pool *pool_of_locked(const void *p)
{
	int i;

	for (i = 0; i < MAX_POOLS; i++)
		if (g_pools[i] && (const uint8_t *)p >= g_pools[i]->base &&
		    (const uint8_t *)p < g_pools[i]->base + (g_pools[i]->npages << PAGE_SHIFT))
			return g_pools[i];
	return NULL;
}

int lz4e_chunk_setup(lz4e_b200_ctx *ctx, size_t pool_pages)
{
	int rc;

	pthread_mutex_lock(&p_mu);
	if (!ctx)
		ctx = current_ctx_locked();
	rc = !ctx ? -ENODEV : pool_get_locked(ctx, pool_pages) ? 0 : -ENOMEM;
	if (!rc && !g_first_ctx)
		g_first_ctx = ctx;
	pthread_mutex_unlock(&p_mu);
	return rc;
}

void lz4e_chunk_bind(lz4e_b200_ctx *ctx)
{
	t_ctx = ctx;
}

void lz4e_chunk_teardown(void)
{
	int i;

	pthread_mutex_lock(&p_mu);
	for (i = 0; i < MAX_POOLS; i++)
		if (g_pools[i] && g_pools[i]->live == 0)
			pool_release_locked(i);
	pthread_mutex_unlock(&p_mu);
}

/* `bytes` of zeroed, page-aligned memory; from the pool when it has room, else
 * plain memory (the codec then bounces it, slower but correct) */
static void *pool_alloc(size_t bytes, int *from_pool)
{
	size_t pages = (bytes + PAGE_SIZE - 1) >> PAGE_SHIFT;
	struct pool *pl;
	void *p = NULL;

	if (!pages)
		pages = 1;
	pthread_mutex_lock(&p_mu);
	pl = pool_get_locked(current_ctx_locked(), 0);
	if (pl && pages < POOL_CLASSES) {
		if (pl->freel[pages]) {
			const size_t idx = pl->freel[pages] - 1;

			p = pl->base + (idx << PAGE_SHIFT);
			pl->freel[pages] = pl->next[idx];
		} else if (pl->bump + pages <= pl->npages) {
			p = pl->base + (pl->bump << PAGE_SHIFT);
			pl->bump += pages;
		}
		if (p)
			pl->live++;
	}
	pthread_mutex_unlock(&p_mu);
	*from_pool = p != NULL;
	if (p)
		return p; /* recycled pinned pages are not cleared: every byte a caller
			   * may read is written first (copy_from_bio / the codec) */
	if (posix_memalign(&p, PAGE_SIZE, pages << PAGE_SHIFT))
		return NULL;
	memset(p, 0, bytes); /* kzalloc in the reference, lz4e_chunk.c:72,79 */
	return p;
}

static void pool_free(void *p, size_t bytes)
{
	size_t pages = (bytes + PAGE_SIZE - 1) >> PAGE_SHIFT;
	struct pool *pl;

	if (!p)
		return;
	if (!pages)
		pages = 1;
	pthread_mutex_lock(&p_mu);
	pl = pool_of_locked(p);
	if (pl) {
		const size_t idx = (size_t)((uint8_t *)p - pl->base) >> PAGE_SHIFT;

		pl->next[idx] = pl->freel[pages];
		pl->freel[pages] = idx + 1;
		pl->live--;
		pthread_mutex_unlock(&p_mu);
		return;
	}
	pthread_mutex_unlock(&p_mu);
	free(p);
}

/*
 * Context a batch of chunks runs on, how many of its buffers are pool buffers
 * of that context and how many pool buffers of it are live: the context of the
 * pool that holds the first chunk's buffers, else this thread's.
 */
static lz4e_b200_ctx *batch_ctx(struct lz4e_chunk **chunks, size_t n, size_t *mine, size_t *live)
{
	struct pool *pl = NULL;
	lz4e_b200_ctx *ctx;
	size_t i;

	*mine = *live = 0;
	pthread_mutex_lock(&p_mu);
	if (n)
		pl = pool_of_locked(chunks[0]->src_buf.data);
	if (!pl && n)
		pl = pool_of_locked(chunks[0]->dst_buf.data);
	if (!pl)
		pl = pool_get_locked(current_ctx_locked(), 0);
	ctx = pl ? pl->ctx : current_ctx_locked();
	if (pl) {
		for (i = 0; i < n; i++)
			*mine += (size_t)(pool_of_locked(chunks[i]->src_buf.data) == pl) +
				 (size_t)(pool_of_locked(chunks[i]->dst_buf.data) == pl);
		*live = pl->live;
	}
	pthread_mutex_unlock(&p_mu);
	return ctx;
}

/* ---- chunks (lz4e_chunk.c:19-107 of the reference) -------------------------- */

void lz4e_buf_copy_from_bio(struct lz4e_buffer *dst, struct bio *src)
{
	char *ptr = dst->data;
	struct bvec_iter it = src->bi_iter;

	while (it.bi_size) {
		const struct bio_vec *v = &src->bi_io_vec[it.bi_idx];
		unsigned int take = v->bv_len - it.bi_bvec_done;

		if (take > it.bi_size)
			take = it.bi_size;
		memcpy(ptr, (char *)page_address(v->bv_page) + v->bv_offset + it.bi_bvec_done,
		       take);
		ptr += take;
		lz4e_bvec_iter_advance(src->bi_io_vec, &it, take);
	}
	dst->data_size = (int)src->bi_iter.bi_size;
}

void lz4e_chunk_free(struct lz4e_chunk *chunk)
{
	if (!chunk)
		return;
	pool_free(chunk->src_buf.data, (size_t)chunk->src_buf.buf_size);
	pool_free(chunk->dst_buf.data, (size_t)chunk->dst_buf.buf_size);
	free(chunk->wrkmem);
	free(chunk);
}

struct lz4e_chunk *lz4e_chunk_alloc(int src_size)
{
	const int dst_size = LZ4E_COMPRESSBOUND(src_size);
	struct lz4e_chunk *chunk;
	int fp;

	if (src_size < 0)
		return NULL;
	chunk = calloc(1, sizeof(*chunk));
	if (!chunk)
		return NULL;
	chunk->src_buf.buf_size = src_size;
	chunk->src_buf.data = pool_alloc((size_t)src_size, &fp);
	if (!chunk->src_buf.data)
		goto free_chunk;
	chunk->dst_buf.buf_size = dst_size;
	chunk->dst_buf.data = pool_alloc((size_t)dst_size, &fp);
	if (!chunk->dst_buf.data)
		goto free_chunk;
	/* the GPU keeps its hash table on chip; wrkmem stays part of the chunk
	 * because LZ4E_compress_default's signature carries it
	 * (the callee clears it, lz4e_compress.c:548; not touched here, so the
	 * allocation costs no page faults) */
	chunk->wrkmem = malloc(LZ4E_MEM_COMPRESS);
	if (!chunk->wrkmem)
		goto free_chunk;
	return chunk;

free_chunk:
	lz4e_chunk_free(chunk);
	return NULL;
}

int lz4e_add_buf_to_bio(struct bio *bio, struct lz4e_buffer *buf)
{
	char *data = buf->data;
	unsigned int left = (unsigned int)buf->buf_size;
	unsigned int off = (unsigned int)((uintptr_t)data & (PAGE_SIZE - 1));
	unsigned int len = left < PAGE_SIZE - off ? left : (unsigned int)(PAGE_SIZE - off);

	while (left) {
		struct bio_vec *v;

		if (bio->bi_vcnt >= bio->bi_max_vecs)
			return -EAGAIN;
		v = &bio->bi_io_vec[bio->bi_vcnt++];
		v->bv_page = (struct page *)((uintptr_t)data & ~(uintptr_t)(PAGE_SIZE - 1));
		v->bv_len = len;
		v->bv_offset = off;
		bio->bi_iter.bi_size += len;
		data += len;
		left -= len;
		off = 0;
		len = left < PAGE_SIZE ? left : (unsigned int)PAGE_SIZE;
	}
	return 0;
}

/* ---- codec calls, batched (lz4e_chunk.c:109-159 of the reference) ----------- */

/* describe a contiguous buffer as one (multi-page) bio_vec */
static void buf_to_bvec(const void *p, unsigned int len, struct bio_vec *bv,
			struct bvec_iter *it)
{
	const uintptr_t a = (uintptr_t)p;

	bv->bv_page = (struct page *)(a & ~(uintptr_t)(PAGE_SIZE - 1));
	bv->bv_offset = (unsigned int)(a & (PAGE_SIZE - 1));
	bv->bv_len = len ? len : 1;
	memset(it, 0, sizeof(*it));
	it->bi_size = len;
}

struct call_ws {
	struct lz4e_b200_req *reqs;
	struct bio_vec *bv;	/* 2 per chunk, for the contiguous forms */
	struct bvec_iter *it;	/* 2 per chunk: throw-away copies, as
				 * lz4e_chunk.c:143-144 makes */
};

static int call_ws_alloc(struct call_ws *w, size_t n)
{
	w->reqs = calloc(n, sizeof(*w->reqs));
	w->bv = calloc(2 * n, sizeof(*w->bv));
	w->it = calloc(2 * n, sizeof(*w->it));
	return (w->reqs && w->bv && w->it) ? 0 : -ENOMEM;
}

static void call_ws_free(struct call_ws *w)
{
	free(w->reqs);
	free(w->bv);
	free(w->it);
}

enum { OP_COMPRESS, OP_COMPRESS_EXT, OP_DECOMPRESS };

static int chunk_batch(struct lz4e_chunk **chunks, size_t n, int *status, int op,
		       unsigned int extra_flags)
{
	struct call_ws w;
	size_t mine = 0, live = 0;
	lz4e_b200_ctx *ctx = batch_ctx(chunks, n, &mine, &live);
	size_t i;
	int rc;

	if (!n)
		return 0;
	if (call_ws_alloc(&w, n)) {
		call_ws_free(&w);
		for (i = 0; i < n; i++)
			status[i] = -ENOMEM;
		return -ENOMEM;
	}
	for (i = 0; i < n; i++) {
		struct lz4e_chunk *c = chunks[i];
		struct lz4e_b200_req *q = &w.reqs[i];

		q->src_iter = &w.it[2 * i];
		q->dst_iter = &w.it[2 * i + 1];
		switch (op) {
		case OP_COMPRESS_EXT:
			q->src = c->src_buf.bio->bi_io_vec;
			q->dst = c->dst_buf.bio->bi_io_vec;
			*q->src_iter = c->src_buf.bio->bi_iter;
			*q->dst_iter = c->dst_buf.bio->bi_iter;
			break;
		case OP_COMPRESS:
			buf_to_bvec(c->src_buf.data, (unsigned int)c->src_buf.data_size,
				    &w.bv[2 * i], q->src_iter);
			buf_to_bvec(c->dst_buf.data, (unsigned int)c->dst_buf.buf_size,
				    &w.bv[2 * i + 1], q->dst_iter);
			q->src = &w.bv[2 * i];
			q->dst = &w.bv[2 * i + 1];
			break;
		default: /* dst_buf.data -> src_buf.data, lz4e_chunk.c:119-137 */
			buf_to_bvec(c->dst_buf.data, (unsigned int)c->dst_buf.data_size,
				    &w.bv[2 * i], q->src_iter);
			buf_to_bvec(c->src_buf.data, (unsigned int)c->src_buf.buf_size,
				    &w.bv[2 * i + 1], q->dst_iter);
			q->src = &w.bv[2 * i];
			q->dst = &w.bv[2 * i + 1];
			break;
		}
	}
	{
		/* The library brings an arena's touched range back in one bulk
		 * copy (lz4e_b200.h).  When pool buffers OUTSIDE this batch are
		 * live (another chunk filled by lz4e_buf_copy_from_bio, say), that
		 * copy must not replace them with the device mirror's stale bytes:
		 * ask for the range to be uploaded first.  A batch that owns every
		 * live pool buffer (the request path below) skips that. */
		unsigned int flags = extra_flags;

		if (live > mine)
			flags |= LZ4E_B200_F_DST_PRESERVE;
		if (op == OP_DECOMPRESS)
			rc = lz4e_b200_decompress_batch(ctx, w.reqs, n, flags);
		else
			rc = lz4e_b200_compress_batch(ctx, w.reqs, n,
						      flags | LZ4E_B200_F_KEEP_ITERS);
	}
	for (i = 0; i < n; i++) {
		struct lz4e_chunk *c = chunks[i];
		const int ret = w.reqs[i].ret;

		if (op == OP_DECOMPRESS) {
			if (rc || ret < 0) { /* lz4e_chunk.c:127-130 */
				status[i] = -EIO;
				continue;
			}
			if (ret != c->src_buf.buf_size) { /* BUG_ON, lz4e_chunk.c:132 */
				fprintf(stderr,
					"lz4e_chunk: decompressed %d bytes, expected %d\n", ret,
					c->src_buf.buf_size);
				abort();
			}
			c->src_buf.data_size = ret;
			status[i] = 0;
		} else {
			if (rc || !ret) { /* lz4e_chunk.c:114-117, :151-154 */
				status[i] = -EIO;
				continue;
			}
			c->dst_buf.data_size = ret;
			status[i] = 0;
		}
	}
	call_ws_free(&w);
	return rc ? -EIO : 0;
}

int lz4e_chunk_compress_batch(struct lz4e_chunk **chunks, size_t n, int *status)
{
	return chunk_batch(chunks, n, status, OP_COMPRESS, 0);
}

int lz4e_chunk_compress_ext_batch(struct lz4e_chunk **chunks, size_t n, int *status)
{
	return chunk_batch(chunks, n, status, OP_COMPRESS_EXT, 0);
}

int lz4e_chunk_decompress_batch(struct lz4e_chunk **chunks, size_t n, int *status)
{
	return chunk_batch(chunks, n, status, OP_DECOMPRESS, 0);
}

int lz4e_chunk_compress(struct lz4e_chunk *chunk)
{
	int st = -EIO;

	chunk_batch(&chunk, 1, &st, OP_COMPRESS, 0);
	return st;
}

int lz4e_chunk_compress_ext(struct lz4e_chunk *chunk)
{
	int st = -EIO;

	chunk_batch(&chunk, 1, &st, OP_COMPRESS_EXT, 0);
	return st;
}

int lz4e_chunk_decompress(struct lz4e_chunk *chunk)
{
	int st = -EIO;

	chunk_batch(&chunk, 1, &st, OP_DECOMPRESS, 0);
	return st;
}

/* ---- request path (lz4e_req.c:53-213, lz4e_stats.c:39-52 of the reference) -- */

struct lz4e_dev {
	uint8_t *store; /* the RAM "underlying device" */
	uint64_t capacity;
	struct lz4e_stats read_stats, write_stats;
	uint64_t compressed_bytes;
	uint64_t compressed_digest; /* sum of FNV-1a-64 over each bio's compressed bytes */
	pthread_mutex_t mu;
};

struct lz4e_dev *lz4e_dev_create(uint64_t capacity)
{
	struct lz4e_dev *dev = calloc(1, sizeof(*dev));

	if (!dev)
		return NULL;
	dev->store = calloc(1, capacity ? capacity : 1);
	if (!dev->store) {
		free(dev);
		return NULL;
	}
	dev->capacity = capacity;
	/* touch every page now: the stand-in device must not take page faults
	 * inside the timed request path */
	memset(dev->store, 0, capacity ? capacity : 1);
	pthread_mutex_init(&dev->mu, NULL);
	return dev;
}

void lz4e_dev_destroy(struct lz4e_dev *dev)
{
	if (!dev)
		return;
	free(dev->store);
	pthread_mutex_destroy(&dev->mu);
	free(dev);
}

/* lz4e_stats_update, lz4e_stats.c:39-52 */
static void stats_update(struct lz4e_stats *st, int failed, unsigned int vcnt,
			 unsigned int bytes)
{
	st->reqs_total++;
	if (failed) {
		st->reqs_failed++;
		return;
	}
	st->vec_count += vcnt;
	st->data_in_bytes += bytes;
}

void lz4e_dev_stats(struct lz4e_dev *dev, int write, struct lz4e_stats *out)
{
	pthread_mutex_lock(&dev->mu);
	*out = write ? dev->write_stats : dev->read_stats;
	pthread_mutex_unlock(&dev->mu);
}

void lz4e_dev_stats_reset(struct lz4e_dev *dev)
{
	pthread_mutex_lock(&dev->mu);
	memset(&dev->read_stats, 0, sizeof(dev->read_stats));
	memset(&dev->write_stats, 0, sizeof(dev->write_stats));
	dev->compressed_bytes = 0;
	dev->compressed_digest = 0;
	pthread_mutex_unlock(&dev->mu);
}

uint64_t lz4e_dev_compressed_bytes(struct lz4e_dev *dev)
{
	return dev->compressed_bytes;
}

uint64_t lz4e_dev_compressed_digest(struct lz4e_dev *dev)
{
	return dev->compressed_digest;
}

/* FNV-1a style digest over 64-bit little-endian words (tail bytes one by one):
 * eight times fewer multiplies than the byte-wise form, which cost more than
 * the codec in the request path */
uint64_t lz4e_fnv1a64(const void *p, size_t n)
{
	const uint8_t *b = p;
	uint64_t h = 1469598103934665603ull ^ (uint64_t)n;
	size_t i = 0;

	for (; i + 8 <= n; i += 8) {
		uint64_t w;

		memcpy(&w, b + i, 8);
		h = (h ^ w) * 1099511628211ull;
		h ^= h >> 29;
	}
	for (; i < n; i++)
		h = (h ^ b[i]) * 1099511628211ull;
	return h;
}

/* lz4e_bio_bytes_to_pages, lz4e_req.c:77-82 */
static unsigned short bytes_to_pages(size_t bytes)
{
	size_t pages = (bytes + PAGE_SIZE - 1) / PAGE_SIZE;

	return (unsigned short)(pages < BIO_MAX_VECS ? pages : BIO_MAX_VECS);
}

static double chunk_now_ms(void)
{
	struct timespec ts;

	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

#define STORE_THREADS 4

struct store_job {
	struct lz4e_dev *dev;
	struct bio **bios;
	struct lz4e_chunk **live;
	const size_t *map;
	const int *st;
	int *status;
	size_t nlive, first, unused, stride;
	uint64_t cbytes, digest;
};

/*
 * Thread `first` of `stride`: the part of every chunk's decompressed bytes that
 * falls into this thread's slice of the RAM device, chunks in submission order
 * (writes of one batch may overlap: the later bio wins, as it does when they
 * are applied one after the other); size and digest of the compressed bytes
 * for every stride-th chunk.
 */
static void *store_main(void *p)
{
	struct store_job *j = p;
	const uint64_t cap = j->dev->capacity;
	const uint64_t lo = cap / j->stride * j->first;
	const uint64_t hi = j->first + 1 == j->stride ? cap : cap / j->stride * (j->first + 1);
	size_t k;

	for (k = 0; k < j->nlive; k++) {
		struct bio *ob = j->bios[j->map[k]];
		struct lz4e_chunk *c = j->live[k];
		const uint64_t pos = (uint64_t)ob->bi_iter.bi_sector * 512u;
		const uint64_t end = pos + (uint64_t)c->src_buf.data_size;
		const uint64_t a = pos > lo ? pos : lo, b = end < hi ? end : hi;

		if (j->st[k]) {
			if (j->first == 0)
				j->status[j->map[k]] = j->st[k];
			continue;
		}
		if (a < b)
			memcpy(j->dev->store + a, (const char *)c->src_buf.data + (a - pos),
			       (size_t)(b - a));
		if (k % j->stride == j->first) {
			j->cbytes += (uint64_t)c->dst_buf.data_size;
			j->digest += lz4e_fnv1a64(c->dst_buf.data, (size_t)c->dst_buf.data_size);
		}
	}
	return NULL;
}

struct wreq {
	struct lz4e_chunk *chunk;
	struct bio new_bio;
	struct bio_vec *vecs;
};

int lz4e_dev_submit_writes(struct lz4e_dev *dev, struct bio **bios, size_t n, int *status)
{
	struct wreq *w = calloc(n ? n : 1, sizeof(*w));
	struct lz4e_chunk **live = calloc(n ? n : 1, sizeof(*live));
	size_t *map = calloc(n ? n : 1, sizeof(*map));
	int *st = calloc(n ? n : 1, sizeof(*st));
	size_t i, k, nlive = 0;
	const int trace = getenv("LZ4E_CHUNK_TRACE") != NULL;
	double t0 = chunk_now_ms(), t1, t2, t3, t4;

	if (!w || !live || !map || !st) {
		for (i = 0; i < n; i++)
			status[i] = -ENOMEM;
		goto out;
	}
	/* lz4e_write_req_init up to the codec call, lz4e_req.c:153-175 */
	for (i = 0; i < n; i++) {
		struct bio *ob = bios[i];
		const unsigned int size = ob->bi_iter.bi_size;
		const uint64_t pos = (uint64_t)ob->bi_iter.bi_sector * 512u;
		unsigned short nv;

		status[i] = 0;
		if (pos + size > dev->capacity || pos + size < pos) {
			status[i] = -ERANGE;
			continue;
		}
		w[i].chunk = lz4e_chunk_alloc((int)size);
		if (!w[i].chunk) {
			status[i] = -ENOMEM;
			continue;
		}
		nv = bytes_to_pages((size_t)w[i].chunk->dst_buf.buf_size + PAGE_SIZE);
		w[i].vecs = calloc(nv ? nv : 1, sizeof(*w[i].vecs));
		if (!w[i].vecs) {
			status[i] = -ENOMEM;
			continue;
		}
		w[i].new_bio.bi_io_vec = w[i].vecs;
		w[i].new_bio.bi_max_vecs = nv;
		w[i].new_bio.bi_opf = ob->bi_opf;
		w[i].new_bio.bi_iter.bi_sector = ob->bi_iter.bi_sector;
		if (lz4e_add_buf_to_bio(&w[i].new_bio, &w[i].chunk->dst_buf)) {
			status[i] = -EIO; /* BLK_STS_IOERR, lz4e_req.c:169-173 */
			continue;
		}
		w[i].chunk->src_buf.bio = ob;
		w[i].chunk->dst_buf.bio = &w[i].new_bio;
		map[nlive] = i;
		live[nlive++] = w[i].chunk;
	}
	/* compress every bio as one block, then decompress it again
	 * (lz4e_req.c:177-189), each step ONE GPU batch */
	t1 = chunk_now_ms();
	lz4e_chunk_compress_ext_batch(live, nlive, st);
	t2 = chunk_now_ms();
	for (k = 0, i = 0; k < nlive; k++) {
		if (st[k]) {
			status[map[k]] = st[k];
			continue;
		}
		map[i] = map[k];
		live[i++] = live[k];
	}
	nlive = i;
	/* the compressed bytes are still current in the device mirror */
	chunk_batch(live, nlive, st, OP_DECOMPRESS, LZ4E_B200_F_SRC_RESIDENT);
	t3 = chunk_now_ms();
	pthread_mutex_lock(&dev->mu);
	{
		/* the new bio is reset and loaded with the DECOMPRESSED buffer
		 * (lz4e_req.c:191-199); the RAM device takes its bytes.  A handful of
		 * threads for a large batch: the copy out of just-DMA-written pinned
		 * pages is the one part of the request path that is pure host memory
		 * traffic (2.5 ms of a 9 ms batch on one thread) */
		struct store_job job = { dev, bios, live, map, st, status, nlive, 0, 0, 1, 0, 0 };
		pthread_t th[STORE_THREADS - 1];
		struct store_job jobs[STORE_THREADS];
		size_t bytes = 0;
		int nt = 1, t;

		for (k = 0; k < nlive; k++)
			bytes += (size_t)live[k]->src_buf.buf_size;
		if (bytes >= ((size_t)2 << 20) && nlive >= 2 * STORE_THREADS)
			nt = STORE_THREADS;
		for (t = 0; t < nt; t++) {
			jobs[t] = job;
			jobs[t].first = (size_t)t;
			jobs[t].stride = (size_t)nt;
		}
		for (t = 1; t < nt; t++)
			if (pthread_create(&th[t - 1], NULL, store_main, &jobs[t])) {
				jobs[t].stride = 0; /* not started: done below */
			}
		store_main(&jobs[0]);
		for (t = 1; t < nt; t++) {
			if (jobs[t].stride)
				pthread_join(th[t - 1], NULL);
			else {
				jobs[t].stride = (size_t)nt;
				store_main(&jobs[t]);
			}
		}
		for (t = 0; t < nt; t++) {
			dev->compressed_bytes += jobs[t].cbytes;
			dev->compressed_digest += jobs[t].digest;
		}
	}
	for (i = 0; i < n; i++)
		stats_update(&dev->write_stats, status[i] != 0,
			     bytes_to_pages(bios[i]->bi_iter.bi_size), bios[i]->bi_iter.bi_size);
	pthread_mutex_unlock(&dev->mu);
	t4 = chunk_now_ms();
	if (trace)
		fprintf(stderr,
			"lz4e_chunk: %zu writes: set-up %.2f ms, compress batch %.2f, verify "
			"(decompress) batch %.2f, store + stats %.2f\n",
			n, t1 - t0, t2 - t1, t3 - t2, t4 - t3);
out:
	for (i = 0; w && i < n; i++) {
		lz4e_chunk_free(w[i].chunk);
		free(w[i].vecs);
	}
	free(w);
	free(live);
	free(map);
	free(st);
	return 0;
}

int lz4e_dev_submit_reads(struct lz4e_dev *dev, struct bio **bios, size_t n, int *status)
{
	size_t i;

	pthread_mutex_lock(&dev->mu);
	for (i = 0; i < n; i++) {
		struct bio *b = bios[i];
		struct bvec_iter it = b->bi_iter;
		const uint64_t pos = (uint64_t)it.bi_sector * 512u;
		const uint8_t *from = dev->store + pos;

		status[i] = 0;
		if (pos + it.bi_size > dev->capacity || pos + it.bi_size < pos) {
			status[i] = -ERANGE;
			stats_update(&dev->read_stats, 1, 0, 0);
			continue;
		}
		while (it.bi_size) {
			const struct bio_vec *v = &b->bi_io_vec[it.bi_idx];
			unsigned int take = v->bv_len - it.bi_bvec_done;

			if (take > it.bi_size)
				take = it.bi_size;
			memcpy((char *)page_address(v->bv_page) + v->bv_offset + it.bi_bvec_done,
			       from, take);
			from += take;
			lz4e_bvec_iter_advance(b->bi_io_vec, &it, take);
		}
		stats_update(&dev->read_stats, 0, b->bi_vcnt, b->bi_iter.bi_size);
	}
	pthread_mutex_unlock(&dev->mu);
	return 0;
}
