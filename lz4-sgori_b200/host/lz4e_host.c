/*
 * lz4e_host.c — the C host library: the reference's LZ4E API shape on top of
 * the CUDA C ABI (include/lz4e_b200.h).
 *
 *   LZ4E_compress_default   ref lz4e/lz4e_compress.c:563-569
 *   LZ4E_decompress_safe    ref lz4e/lz4e_decompress.c:462-470
 *   lz4e_b200_*_batch       arrays of the same argument tuple, the shape the
 *                           chunk layer (lz4e_bdev/lz4e_chunk.c:139-159) uses
 *
 * What happens here is only what the reference's iterator helpers do on the
 * CPU side of the boundary: walk bio_vec arrays under a bvec_iter
 * (LZ4E_for_each_bvec, lz4e/include/lz4e_defs.h:118-123), resolve pages,
 * and move iterators (bvec_iter_advance).  All codec work is enqueued on the
 * GPU; without a device every entry point fails loudly.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/lz4e.h"
#include "../../include/lz4e_b200.h"
#include "lz4e_internal.h"

#define MAX_ARENAS 16
#define NLANES 3 /* chunks in flight: upload / kernel / download overlap */

/* snapshot of the context's arenas, taken once per batch */
struct arena_set {
	struct {
		lz4e_b200_arena *a;
		uint8_t *host, *host_end, *dev;
	} v[MAX_ARENAS];
	int n, last;
};

struct range {
	size_t lo, hi; /* byte range touched inside the arena; lo == hi: none */
};

/* a host segment that is not inside any arena: copied through device bounce
 * memory */
struct bounce {
	uint8_t *host;
	size_t dev_off; /* offset inside the bounce buffer */
	uint32_t len;
	uint32_t blk;
	uint32_t pos; /* linear position of the segment's first byte in the block */
};

struct side {
	struct lz4e_b200_dev_bvec *bv; /* in the pinned workspace */
	struct lz4e_b200_dev_iter *it;
	size_t nbv;
	struct range ar[MAX_ARENAS];
	size_t sum[MAX_ARENAS]; /* bytes of this side's segments inside each arena */
	struct bounce *bn;
	size_t nbn, cap_bn;
	size_t bounce_bytes;
	uint32_t max_size;
	int stage_all;	/* pack every segment (arena pages too) into the staging buffer */
	int all_linear; /* every block is a single segment */
	int any_u16;	/* some block would select byU16 (compress source) */
	/* regular layout: every block is ONE segment inside the same arena and the
	 * blocks' first bytes are reg_pitch bytes apart (an array of slots) */
	int reg_ok, reg_arena;
	size_t reg_first, reg_prev, reg_pitch;
};

/*
 * One chunk of a batch in flight.  A batch is cut into chunks of about
 * LZ4E_B200_CHUNK_MB (default 32) MiB; chunk c runs on lane c % NLANES, each
 * lane with its own stream and workspace, so the page upload of chunk c+1 and
 * the download of chunk c-1 overlap the kernel of chunk c.  Kernels are
 * chained by an event and run one after the other (each one fills the GPU, and
 * the decompressor's scratch is per context).
 */
struct lane {
	void *stream, *ev_kernel;
	void *ev_t[4]; /* LZ4E_B200_TRACE=2: start, uploads done, kernel done, results down */
	double t_submit;
	size_t chunk_no, chunk_bytes;
	uint8_t *pin, *dev, *bounce;
	uint8_t *stage; /* pinned image of the source part of `bounce` */
	size_t pin_bytes, bounce_bytes, stage_bytes;
	int busy;
	struct lz4e_b200_req *reqs;
	size_t n;
	struct side src, dst;
	size_t o_ret, o_end;
};

/*
 * One workspace per CONTEXT (attached to it, released by
 * lz4e_b200_ctx_destroy): batches on different GPUs of one process run side by
 * side, batches on the same context take turns (SURVEY.md section 8b
 * "Threading": callers in many task contexts share a lazily created pool).
 */
#define MAX_WORKERS 16

/* a few helper threads per context for the one host-side job that is pure
 * memory traffic: packing scattered source bytes into a lane's pinned staging
 * buffer (pack_run) */
struct workers {
	pthread_t th[MAX_WORKERS];
	pthread_mutex_t mu;
	pthread_cond_t go, done;
	int n, started, quit;
	unsigned long gen;
	int pending;
	/* the job: entries [0, count) of bn, split in n + 1 contiguous parts */
	const struct bounce *bn;
	size_t count;
	uint8_t *base;
};

struct host_ws {
	pthread_mutex_t mu;
	int ready;
	struct workers wk;
	struct lane lane[NLANES];
	void *ev_batch; /* LZ4E_B200_TRACE=2: start of the batch on the device */
	int tracing;
	double t_batch;
	/* LZ4E_B200_TRACE=1: where a batch spends its host time (stderr) */
	struct {
		double build, enqueue, wait, pack;
		size_t bulk_up, bulk_down, kern_up, kern_down, packed_up, rows_up; /* chunks per mode */
	} trace;
};

static pthread_mutex_t g_init_mu = PTHREAD_MUTEX_INITIALIZER;

static void pack_part(const struct bounce *bn, size_t count, uint8_t *base, int part, int parts)
{
	size_t lo = count * (size_t)part / (size_t)parts;
	const size_t hi = count * ((size_t)part + 1) / (size_t)parts;

	for (; lo < hi; lo++)
		memcpy(base + bn[lo].dev_off, bn[lo].host, bn[lo].len);
}

struct worker_arg {
	struct workers *wk;
	int idx;
};

static void *worker_main(void *p)
{
	struct worker_arg *wa = p;
	struct workers *wk = wa->wk;
	const int idx = wa->idx;
	unsigned long seen = 0;

	free(wa);
	pthread_mutex_lock(&wk->mu);
	for (;;) {
		while (!wk->quit && wk->gen == seen)
			pthread_cond_wait(&wk->go, &wk->mu);
		if (wk->quit)
			break;
		seen = wk->gen;
		pthread_mutex_unlock(&wk->mu);
		pack_part(wk->bn, wk->count, wk->base, idx + 1, wk->n + 1);
		pthread_mutex_lock(&wk->mu);
		if (--wk->pending == 0)
			pthread_cond_signal(&wk->done);
	}
	pthread_mutex_unlock(&wk->mu);
	return NULL;
}

static void workers_start(struct workers *wk)
{
	const char *e = getenv("LZ4E_B200_HOST_THREADS");
	int want = e ? atoi(e) : 4, i;

	pthread_mutex_init(&wk->mu, NULL);
	pthread_cond_init(&wk->go, NULL);
	pthread_cond_init(&wk->done, NULL);
	wk->started = 1;
	if (want > MAX_WORKERS + 1)
		want = MAX_WORKERS + 1;
	for (i = 0; i + 1 < want; i++) { /* the calling thread is one of them */
		struct worker_arg *wa = malloc(sizeof(*wa));

		if (!wa)
			break;
		wa->wk = wk;
		wa->idx = wk->n;
		if (pthread_create(&wk->th[wk->n], NULL, worker_main, wa)) {
			free(wa);
			break;
		}
		wk->n++;
	}
}

static void workers_stop(struct workers *wk)
{
	int i;

	if (!wk->started)
		return;
	pthread_mutex_lock(&wk->mu);
	wk->quit = 1;
	pthread_cond_broadcast(&wk->go);
	pthread_mutex_unlock(&wk->mu);
	for (i = 0; i < wk->n; i++)
		pthread_join(wk->th[i], NULL);
	pthread_cond_destroy(&wk->go);
	pthread_cond_destroy(&wk->done);
	pthread_mutex_destroy(&wk->mu);
	wk->started = 0;
}

/* copy every bn[i] (host bytes) to base + bn[i].dev_off, with the helpers when
 * the job is large enough to pay for waking them */
static void pack_run(struct workers *wk, const struct bounce *bn, size_t count, uint8_t *base,
		     size_t bytes)
{
	if (!count)
		return;
	if (wk->n == 0 || bytes < ((size_t)1 << 20)) {
		pack_part(bn, count, base, 0, 1);
		return;
	}
	pthread_mutex_lock(&wk->mu);
	wk->bn = bn;
	wk->count = count;
	wk->base = base;
	wk->pending = wk->n;
	wk->gen++;
	pthread_cond_broadcast(&wk->go);
	pthread_mutex_unlock(&wk->mu);
	pack_part(bn, count, base, 0, wk->n + 1);
	pthread_mutex_lock(&wk->mu);
	while (wk->pending)
		pthread_cond_wait(&wk->done, &wk->mu);
	pthread_mutex_unlock(&wk->mu);
}

#define NWS 2 /* batches in flight per context */

/* what is attached to the context: NWS workspaces, handed out to batches */
struct host_pool {
	struct host_ws ws[NWS];
};

static void ws_release(lz4e_b200_ctx *ctx, struct host_ws *ws)
{
	int l;

	if (!ws->ready)
		return;
	workers_stop(&ws->wk);
	for (l = 0; l < NLANES; l++) {
		struct lane *ln = &ws->lane[l];

		lz4e_b200_host_free(ctx, ln->stage);
		lz4e_b200_host_free(ctx, ln->pin);
		lz4e_b200_dev_free(ctx, ln->dev);
		lz4e_b200_dev_free(ctx, ln->bounce);
		if (ln->ev_kernel)
			lz4e_b200_event_destroy(ctx, ln->ev_kernel);
		for (int k = 0; k < 4; k++)
			if (ln->ev_t[k])
				lz4e_b200_event_destroy(ctx, ln->ev_t[k]);
		if (ln->stream)
			lz4e_b200_stream_destroy(ctx, ln->stream);
		free(ln->src.bn);
		free(ln->dst.bn);
	}
	if (ws->ev_batch)
		lz4e_b200_event_destroy(ctx, ws->ev_batch);
	pthread_mutex_destroy(&ws->mu);
	ws->ready = 0;
}

static void ws_dtor(lz4e_b200_ctx *ctx, void *p)
{
	struct host_pool *hp = p;
	int k;

	for (k = 0; k < NWS; k++)
		ws_release(ctx, &hp->ws[k]);
}

static int ws_init(lz4e_b200_ctx *ctx, struct host_ws *ws)
{
	int l, ok = 1;

	pthread_mutex_lock(&g_init_mu);
	if (!ws->ready) {
		for (l = 0; l < NLANES; l++) {
			ws->lane[l].stream = lz4e_b200_stream_create(ctx);
			ws->lane[l].ev_kernel = lz4e_b200_event_create(ctx);
			if (!ws->lane[l].stream || !ws->lane[l].ev_kernel)
				ok = 0;
		}
		if (ok) {
			pthread_mutex_init(&ws->mu, NULL);
			workers_start(&ws->wk);
			ws->ready = 1;
		}
	}
	pthread_mutex_unlock(&g_init_mu);
	return ws->ready;
}

/*
 * A workspace of the context, LOCKED: the first one that is free, so that two
 * threads (a compress batch and a decompress batch, say: one uploads, the other
 * downloads, and PCIe is full duplex) run side by side on one GPU; a third
 * caller waits for the first workspace.  Created on first use; NULL when that
 * fails.
 */
static struct host_ws *ws_acquire(lz4e_b200_ctx *ctx)
{
	struct host_pool *hp = lz4e_ctx_attach(ctx, LZ4E_ATTACH_HOST_WS, sizeof(*hp), ws_dtor);
	int k;

	if (!hp)
		return NULL;
	for (k = 0; k < NWS; k++) {
		if (!hp->ws[k].ready && !ws_init(ctx, &hp->ws[k]))
			return NULL;
		if (pthread_mutex_trylock(&hp->ws[k].mu) == 0)
			return &hp->ws[k];
	}
	pthread_mutex_lock(&hp->ws[0].mu);
	return &hp->ws[0];
}

/* the lane is idle when this is called */
static int lane_reserve(lz4e_b200_ctx *ctx, struct lane *ln, size_t pin_bytes,
			size_t bounce_bytes, size_t stage_bytes)
{
	if (stage_bytes > ln->stage_bytes) {
		const size_t want = stage_bytes + stage_bytes / 4 + 4096;

		lz4e_b200_host_free(ctx, ln->stage);
		ln->stage = lz4e_b200_host_alloc(ctx, want);
		ln->stage_bytes = ln->stage ? want : 0;
		if (!ln->stage)
			return LZ4E_B200_ENOMEM;
	}
	if (pin_bytes > ln->pin_bytes) {
		const size_t want = pin_bytes + pin_bytes / 4 + 4096;

		lz4e_b200_host_free(ctx, ln->pin);
		lz4e_b200_dev_free(ctx, ln->dev);
		ln->pin = lz4e_b200_host_alloc(ctx, want);
		ln->dev = lz4e_b200_dev_alloc(ctx, want);
		ln->pin_bytes = (ln->pin && ln->dev) ? want : 0;
		if (!ln->pin_bytes)
			return LZ4E_B200_ENOMEM;
	}
	if (bounce_bytes > ln->bounce_bytes) {
		const size_t want = bounce_bytes + bounce_bytes / 4 + 4096;

		lz4e_b200_dev_free(ctx, ln->bounce);
		ln->bounce = lz4e_b200_dev_alloc(ctx, want);
		ln->bounce_bytes = ln->bounce ? want : 0;
		if (!ln->bounce)
			return LZ4E_B200_ENOMEM;
	}
	return LZ4E_B200_OK;
}

static void arena_set_init(lz4e_b200_ctx *ctx, struct arena_set *as)
{
	lz4e_b200_arena *list[MAX_ARENAS];
	size_t i, n = lz4e_b200_arena_snapshot(ctx, list, MAX_ARENAS);

	as->n = (int)n;
	as->last = 0;
	for (i = 0; i < n; i++) {
		as->v[i].a = list[i];
		as->v[i].host = (uint8_t *)lz4e_b200_arena_pages(list[i]);
		as->v[i].host_end = as->v[i].host + lz4e_b200_arena_npages(list[i]) * PAGE_SIZE;
		as->v[i].dev = (uint8_t *)lz4e_b200_arena_dev_base(list[i]);
	}
}

/* index of the arena holding [p, p + len), or -1 */
static int arena_find(struct arena_set *as, const uint8_t *p, size_t len)
{
	int k = as->last, i;

	if (k < as->n && p >= as->v[k].host && p + len <= as->v[k].host_end)
		return k;
	for (i = 0; i < as->n; i++) {
		if (p >= as->v[i].host && p + len <= as->v[i].host_end) {
			as->last = i;
			return i;
		}
	}
	return -1;
}

static void range_add(struct range *r, size_t lo, size_t hi)
{
	if (r->lo == r->hi) {
		r->lo = lo;
		r->hi = hi;
		return;
	}
	if (lo < r->lo)
		r->lo = lo;
	if (hi > r->hi)
		r->hi = hi;
}

static int ranges_overlap(const struct range *a, const struct range *b)
{
	int k;

	for (k = 0; k < MAX_ARENAS; k++)
		if (a[k].lo < a[k].hi && b[k].lo < b[k].hi && a[k].lo < b[k].hi &&
		    b[k].lo < a[k].hi)
			return 1;
	return 0;
}

/* hull and byte count of the host addresses a side's segments cover */
struct span {
	uintptr_t lo, hi;
	uint64_t sum;
};

/* count_bvecs() that also grows the span */
static size_t scan_bvecs(const struct bio_vec *bv, const struct bvec_iter *it, struct span *sp)
{
	size_t n = 0;
	unsigned int left = it->bi_size, idx = it->bi_idx, done = it->bi_bvec_done;

	while (left) {
		const struct bio_vec *v = &bv[idx];
		unsigned int take = v->bv_len - done;
		const uintptr_t a = (uintptr_t)page_address(v->bv_page) + v->bv_offset + done;

		if (take > left)
			take = left;
		if (a < sp->lo)
			sp->lo = a;
		if (a + take > sp->hi)
			sp->hi = a + take;
		sp->sum += take;
		left -= take;
		done = 0;
		idx++;
		n++;
	}
	return n ? n : 1;
}

/* number of bio_vecs under an iterator (LZ4E_for_each_bvec) */
static size_t count_bvecs(const struct bio_vec *bv, const struct bvec_iter *it)
{
	size_t n = 0;
	unsigned int left = it->bi_size, idx = it->bi_idx, done = it->bi_bvec_done;

	while (left) {
		unsigned int take = bv[idx].bv_len - done;

		if (take > left)
			take = left;
		left -= take;
		done = 0;
		idx++;
		n++;
	}
	return n ? n : 1; /* an empty block still gets one (dummy) entry */
}

/* Append the bio_vecs block `blk` covers to side s.  The device iterator is
 * rebased so that its bi_idx is 0 (indices relative to the start are all the
 * reference's table-type rule looks at, lz4e_compress.c:195-202). */
static int side_add(struct arena_set *as, struct side *s, uint32_t blk,
		    const struct bio_vec *bv, const struct bvec_iter *it)
{
	unsigned int left = it->bi_size, idx = it->bi_idx, done = it->bi_bvec_done;
	struct lz4e_b200_dev_iter *di = &s->it[blk];
	uint32_t pos = 0, nseg = 0;
	int u16 = 1;

	di->first = (uint32_t)s->nbv;
	di->bi_size = it->bi_size;
	di->bi_idx = 0;
	di->bi_bvec_done = done;
	if (it->bi_size > s->max_size)
		s->max_size = it->bi_size;
	if (!left) {
		s->bv[s->nbv].addr = 0;
		s->bv[s->nbv].len = 0;
		s->bv[s->nbv].rsvd = 0;
		s->nbv++;
		s->reg_ok = 0;
		return LZ4E_B200_OK;
	}
	while (left) {
		const struct bio_vec *v = &bv[idx];
		unsigned int take = v->bv_len - done;
		uint8_t *h = (uint8_t *)page_address(v->bv_page) + v->bv_offset;
		struct lz4e_b200_dev_bvec *o = &s->bv[s->nbv++];
		int k;

		if (take > left)
			take = left;
		if (nseg >= 16 || v->bv_len > 4096)
			u16 = 0;
		k = s->stage_all ? -1 : arena_find(as, h + done, take);
		if (k >= 0) {
			const size_t off = (size_t)((h + done) - as->v[k].host);

			/* address of the bvec's byte 0 */
			o->addr = (uint64_t)(uintptr_t)(as->v[k].dev + off) - done;
			o->rsvd = 0;
			range_add(&s->ar[k], off, off + take);
			s->sum[k] += take;
			if (s->reg_ok && nseg == 0) {
				if (blk == 0) {
					s->reg_arena = k;
					s->reg_first = off;
				} else if (k != s->reg_arena || off <= s->reg_prev ||
					   (blk > 1 && off - s->reg_prev != s->reg_pitch)) {
					s->reg_ok = 0;
				} else if (blk == 1) {
					s->reg_pitch = off - s->reg_prev;
				}
				s->reg_prev = off;
			}
		} else {
			s->reg_ok = 0;
			/* bounce: keep the host address's alignment mod 16 */
			size_t off = (s->bounce_bytes + 15) & ~(size_t)15;
			struct bounce *b;

			off += (size_t)((uintptr_t)(h + done) & 15);
			if (s->nbn == s->cap_bn) {
				size_t nc = s->cap_bn ? s->cap_bn * 2 : 64;
				struct bounce *nb = realloc(s->bn, nc * sizeof(*nb));

				if (!nb)
					return LZ4E_B200_ENOMEM;
				s->bn = nb;
				s->cap_bn = nc;
			}
			b = &s->bn[s->nbn++];
			b->host = h + done;
			b->dev_off = off;
			b->len = take;
			b->blk = blk;
			b->pos = pos;
			s->bounce_bytes = off + take;
			o->addr = (uint64_t)off - done; /* rebased once allocated */
			o->rsvd = 1;
		}
		o->len = v->bv_len;
		left -= take;
		pos += take;
		done = 0;
		idx++;
		nseg++;
	}
	if (nseg > 1) {
		s->all_linear = 0;
		s->reg_ok = 0;
	}
	if (u16)
		s->any_u16 = 1;
	return LZ4E_B200_OK;
}

static void side_reset(struct side *s)
{
	struct bounce *bn = s->bn;
	const size_t cap = s->cap_bn;

	memset(s, 0, sizeof(*s));
	s->bn = bn;
	s->cap_bn = cap;
	s->all_linear = 1;
	s->reg_ok = 1;
	s->reg_arena = -1;
}

static void side_rebase_bounce(struct side *s, uint8_t *bounce_base)
{
	size_t i;

	for (i = 0; i < s->nbv; i++) {
		if (s->bv[i].rsvd) {
			s->bv[i].addr += (uint64_t)(uintptr_t)bounce_base;
			s->bv[i].rsvd = 0;
		}
	}
}

static double now_ms(void)
{
	struct timespec ts;

	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static size_t al256(size_t x)
{
	return (x + 255) & ~(size_t)255;
}

static int copy_ranges(lz4e_b200_ctx *ctx, struct arena_set *as, void *stream,
		       const struct range *r, int to_device)
{
	int k, rc = LZ4E_B200_OK;

	for (k = 0; k < as->n && !rc; k++) {
		if (r[k].lo == r[k].hi)
			continue;
		if (to_device)
			rc = lz4e_b200_copy_h2d(ctx, stream, as->v[k].dev + r[k].lo,
						as->v[k].host + r[k].lo, r[k].hi - r[k].lo);
		else
			rc = lz4e_b200_copy_d2h(ctx, stream, as->v[k].host + r[k].lo,
						as->v[k].dev + r[k].lo, r[k].hi - r[k].lo);
	}
	return rc;
}

/*
 * A side is DENSE when, in every arena it touches, its segments tile the byte
 * range they span: a copy-engine copy of that range then moves no byte that
 * belongs to somebody else.  Everything else goes through the exact-byte
 * transfer kernel (lz4e_b200_xfer_dev).
 */
static int side_dense(const struct side *s)
{
	int k;

	for (k = 0; k < MAX_ARENAS; k++)
		if (s->ar[k].hi - s->ar[k].lo != s->sum[k])
			return 0;
	return 1;
}

static int side_in_arenas(const struct side *s)
{
	int k;

	for (k = 0; k < MAX_ARENAS; k++)
		if (s->sum[k])
			return 1;
	return 0;
}

/* LZ4E_B200_XFER: 0 = copy engines only (ranges, the round-1 behaviour),
 * 1 = default (kernel for compressed bytes and sparse sides, copy engines for
 * dense uncompressed sides), 2 = transfer kernel for everything */
static int xfer_mode(void)
{
	const char *e = getenv("LZ4E_B200_XFER");
	const int mode = e ? atoi(e) : 1;

	return mode < 0 || mode > 2 ? 1 : mode;
}

/* LZ4E_B200_SPARSE_SRC: how sources that do not tile their address range go
 * up: "kernel" (transfer kernel, exact bytes), "pack" (host threads pack them
 * into pinned staging, one copy-engine transfer), "hull" (copy engine over the
 * whole range they span: also moves the bytes in between) */
enum { SPARSE_KERNEL, SPARSE_PACK, SPARSE_HULL };

static int sparse_src_mode(const struct side *s)
{
	const char *e = getenv("LZ4E_B200_SPARSE_SRC");
	size_t sum = 0, span = 0;
	int k;

	if (e && !strcmp(e, "pack"))
		return SPARSE_PACK;
	if (e && !strcmp(e, "hull"))
		return SPARSE_HULL;
	if (e && !strcmp(e, "kernel"))
		return SPARSE_KERNEL;
	if (!s)
		return SPARSE_KERNEL;
	/*
	 * Measured on B200 / PCIe 5 (profiles/r2k_e2e_timeline.md): SM reads of
	 * host memory stop at about 30 GB/s and queue behind concurrent
	 * device -> host writes (PCIe ordering), the copy engines read at 47-55
	 * GB/s whatever else is going on; host threads pack at 16-20 GB/s.  So a
	 * side that fills a quarter or more of the range it spans (compressed
	 * 4 KiB blocks in their LZ4E_COMPRESSBOUND slots: 40 %) goes up as that
	 * range, sparser ones through the kernel.
	 */
	for (k = 0; k < MAX_ARENAS; k++) {
		sum += s->sum[k];
		span += s->ar[k].hi - s->ar[k].lo;
	}
	return span && sum * 4 >= span ? SPARSE_HULL : SPARSE_KERNEL;
}

/*
 * Regular sparse sources (compressed blocks of very different sizes in equal
 * slots): the first W bytes of every slot go up as the rows of ONE 2-D
 * copy-engine transfer, what a block has beyond W through the transfer kernel.
 * W minimises rows + 2 x excess (SM reads over PCIe cost about twice a copy-
 * engine byte).  Returns W (a multiple of 64), or 0 when the spanned range is
 * not worth cutting.
 */
static uint32_t pick_row_width(const struct side *s, size_t n, uint64_t *excess_out)
{
	uint64_t hist[65], best = 0, above_cnt = 0, above_sum = 0, sum = 0;
	uint32_t bw = 64, best_w = 0;
	size_t i;
	int j;

	if (!s->reg_ok || n < 64 || !s->reg_pitch || s->max_size > s->reg_pitch)
		return 0;
	while ((uint64_t)bw * 64 < s->max_size)
		bw *= 2;
	memset(hist, 0, sizeof(hist));
	for (i = 0; i < n; i++) {
		hist[(s->it[i].bi_size + bw - 1) / bw]++; /* sizes rounded up to a bucket edge */
		sum += s->it[i].bi_size;
	}
	/* walk W down from the top bucket edge; `above` = blocks longer than W */
	for (j = 64; j >= 1; j--) {
		const uint64_t w = (uint64_t)j * bw;
		/* excess over w of the blocks in buckets above j (upper bound) */
		const uint64_t cost = n * w + 2 * (above_sum - above_cnt * w);

		if (w <= s->reg_pitch && (!best_w || cost < best)) {
			best = cost;
			best_w = (uint32_t)w;
			*excess_out = above_sum - above_cnt * w;
		}
		above_cnt += hist[j];
		above_sum += hist[j] * (uint64_t)j * bw;
	}
	/* against the whole range by copy engine (n x pitch) */
	if (!best_w || best * 10 > (uint64_t)n * s->reg_pitch * 8)
		return 0;
	(void)sum;
	return best_w;
}

/* wait for a chunk and hand its results back to the caller's requests */
static int lane_finish(lz4e_b200_ctx *ctx, struct host_ws *ws, struct lane *ln,
		       int decompress, unsigned flags)
{
	const int32_t *rets = (const int32_t *)(ln->pin + ln->o_ret);
	const struct lz4e_b200_dev_end *ends =
		(const struct lz4e_b200_dev_end *)(ln->pin + ln->o_end);
	size_t i;
	int rc;

	double t0;

	if (!ln->busy)
		return LZ4E_B200_OK;
	ln->busy = 0;
	t0 = now_ms();
	rc = lz4e_b200_sync(ctx, ln->stream);
	ws->trace.wait += now_ms() - t0;
	if (ws->tracing >= 2 && !rc && ln->ev_t[0])
		fprintf(stderr,
			"lz4e_b200:   chunk %3zu (%6.1f MiB) host: submitted at %6.2f ms, finished at "
			"%6.2f; device: starts %6.2f, uploads %5.2f, kernel (incl. wait for the one "
			"before) %5.2f, results down %5.2f ms\n",
			ln->chunk_no, ln->chunk_bytes / 1048576.0, ln->t_submit - ws->t_batch,
			now_ms() - ws->t_batch,
			lz4e_b200_event_elapsed_ms(ctx, ws->ev_batch, ln->ev_t[0]),
			lz4e_b200_event_elapsed_ms(ctx, ln->ev_t[0], ln->ev_t[1]),
			lz4e_b200_event_elapsed_ms(ctx, ln->ev_t[1], ln->ev_t[2]),
			lz4e_b200_event_elapsed_ms(ctx, ln->ev_t[2], ln->ev_t[3]));
	/* bounced destination segments: bring back exactly the bytes produced */
	for (i = 0; i < ln->dst.nbn && !rc; i++) {
		const struct bounce *b = &ln->dst.bn[i];
		const int32_t r = rets[b->blk];
		uint32_t len;

		if (r <= 0 || (uint32_t)r <= b->pos)
			continue;
		len = (uint32_t)r - b->pos;
		if (len > b->len)
			len = b->len;
		rc = lz4e_b200_copy_d2h(ctx, ln->stream, b->host, ln->bounce + b->dev_off, len);
	}
	if (!rc && ln->dst.nbn)
		rc = lz4e_b200_sync(ctx, ln->stream);
	if (rc)
		return rc;
	for (i = 0; i < ln->n; i++) {
		struct lz4e_b200_req *q = &ln->reqs[i];

		q->ret = rets[i];
		if (!decompress && rets[i] >= 0 && !(flags & LZ4E_B200_F_KEEP_ITERS)) {
			/* leave the iterators where the reference does,
			 * lz4e_compress.c:304-305 and :528-529 — also after a 0
			 * return (:363, :430, :509), where they stay at the failing
			 * check; blocks rejected up front report 0/0 */
			lz4e_bvec_iter_advance(q->src, q->src_iter, ends[i].src_pos);
			lz4e_bvec_iter_advance(q->dst, q->dst_iter, ends[i].dst_pos);
		}
	}
	return LZ4E_B200_OK;
}

/* build the tables of one chunk in the lane's workspace and enqueue its
 * uploads, kernel and downloads on the lane's stream */
static int lane_submit(lz4e_b200_ctx *ctx, struct host_ws *ws, struct arena_set *as,
		       struct lane *ln, struct lz4e_b200_req *reqs, size_t n, size_t ns,
		       size_t nd, unsigned flags, int decompress, void *prev_kernel,
		       int stage_src)
{
	size_t i, off, pin_bytes, o_sbv, o_dbv, o_sit, o_dit, sb;
	int rc = LZ4E_B200_OK, l, src_kernel, dst_kernel;
	uint32_t row_w = 0;
	uint64_t row_excess = 0;
	const int mode = xfer_mode();
	double t0 = now_ms(), t1;
	struct lz4e_b200_dev_bvec *d_sbv, *d_dbv;
	struct lz4e_b200_dev_iter *d_sit, *d_dit;

	off = 0;
	o_sbv = off; off = al256(off + ns * sizeof(struct lz4e_b200_dev_bvec));
	o_dbv = off; off = al256(off + nd * sizeof(struct lz4e_b200_dev_bvec));
	o_sit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	o_dit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	ln->o_ret = off; off = al256(off + n * sizeof(int32_t));
	ln->o_end = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_end));
	pin_bytes = off;
	rc = lane_reserve(ctx, ln, pin_bytes, 0, 0);
	if (rc)
		return rc;
	ln->reqs = reqs;
	ln->n = n;
	side_reset(&ln->src);
	side_reset(&ln->dst);
	ln->src.stage_all = stage_src;
	ln->src.bv = (struct lz4e_b200_dev_bvec *)(ln->pin + o_sbv);
	ln->dst.bv = (struct lz4e_b200_dev_bvec *)(ln->pin + o_dbv);
	ln->src.it = (struct lz4e_b200_dev_iter *)(ln->pin + o_sit);
	ln->dst.it = (struct lz4e_b200_dev_iter *)(ln->pin + o_dit);
	for (i = 0; i < n && !rc; i++) {
		rc = side_add(as, &ln->src, (uint32_t)i, reqs[i].src, reqs[i].src_iter);
		if (!rc)
			rc = side_add(as, &ln->dst, (uint32_t)i, reqs[i].dst, reqs[i].dst_iter);
	}
	if (rc)
		return rc;
	d_sbv = (struct lz4e_b200_dev_bvec *)(ln->dev + o_sbv);
	d_dbv = (struct lz4e_b200_dev_bvec *)(ln->dev + o_dbv);
	d_sit = (struct lz4e_b200_dev_iter *)(ln->dev + o_sit);
	d_dit = (struct lz4e_b200_dev_iter *)(ln->dev + o_dit);
	/*
	 * How the pages travel.  Source: copy engines when the side is dense
	 * (uncompressed blocks laid end to end), the transfer kernel otherwise —
	 * compressed blocks sitting in LZ4E_COMPRESSBOUND-sized slots, scattered
	 * segments with foreign pages in between.  Destination: a compressor's
	 * output always through the kernel (rets[b] bytes per block, known only on
	 * the device); a decompressor's through the copy engines when dense.
	 * LZ4E_B200_F_DST_PRESERVE asks that no other host byte changes: kernel.
	 */
	src_kernel = mode == 2 || (mode == 1 && !side_dense(&ln->src) &&
				   sparse_src_mode(&ln->src) != SPARSE_HULL);
	dst_kernel = mode == 2 ||
		     (mode == 1 && (!decompress || !side_dense(&ln->dst) ||
				    (flags & LZ4E_B200_F_DST_PRESERVE)));
	if (!side_in_arenas(&ln->src))
		src_kernel = 0;
	/* sparse sources in equal slots as rows + excess: LZ4E_B200_SPARSE_SRC=rows.
	 * Measured (profiles/r2k_e2e_timeline.md): 1.70 instead of 2.21 GB go up
	 * per GiB step, but rows of under 2 KiB keep the copy engine below its
	 * rate and the decompress batch takes 31.1 instead of 29.5 ms: not the
	 * default */
	{
		const char *e = getenv("LZ4E_B200_SPARSE_SRC");

		if (mode == 1 && !side_dense(&ln->src) && e && !strcmp(e, "rows"))
			row_w = pick_row_width(&ln->src, n, &row_excess);
	}
	if (!side_in_arenas(&ln->dst))
		dst_kernel = 0;
	t1 = now_ms();
	ws->trace.build += t1 - t0;
	/*
	 * Chunks still in flight: a copy of this chunk must not cross bytes a
	 * running chunk produces (or the other way round).  The ranges are hulls,
	 * so callers that interleave the pages of different chunks inside one
	 * arena lose the overlap, not the result.
	 */
	for (l = 0; l < NLANES && !rc; l++) {
		struct lane *o = &ws->lane[l];

		if (o == ln || !o->busy)
			continue;
		if (ranges_overlap(ln->src.ar, o->dst.ar) || ranges_overlap(ln->dst.ar, o->dst.ar) ||
		    ranges_overlap(ln->dst.ar, o->src.ar))
			rc = lane_finish(ctx, ws, o, decompress, flags);
	}
	if (rc)
		return rc;
	/* bounce space: source segments first, destination segments after */
	sb = (ln->src.bounce_bytes + 255) & ~(size_t)255;
	if (ln->src.nbn || ln->dst.nbn) {
		rc = lane_reserve(ctx, ln, pin_bytes, sb + ln->dst.bounce_bytes + 256,
				  ln->src.bounce_bytes);
		if (rc)
			return rc;
		side_rebase_bounce(&ln->src, ln->bounce);
		side_rebase_bounce(&ln->dst, ln->bounce + sb);
		if (ln->src.nbn) {
			/* scattered source bytes: packed into pinned memory by the
			 * host threads, then ONE copy-engine transfer */
			const double tp = now_ms();

			pack_run(&ws->wk, ln->src.bn, ln->src.nbn, ln->stage, ln->src.bounce_bytes);
			ws->trace.pack += now_ms() - tp;
			rc = lz4e_b200_copy_h2d(ctx, ln->stream, ln->bounce, ln->stage,
						ln->src.bounce_bytes);
			ws->trace.packed_up++;
		}
		for (i = 0; i < ln->dst.nbn; i++)
			ln->dst.bn[i].dev_off += sb;
	}
	ln->t_submit = now_ms();
	if (ws->tracing >= 2) {
		int k;

		for (k = 0; k < 4; k++)
			if (!ln->ev_t[k])
				ln->ev_t[k] = lz4e_b200_event_create_timed(ctx);
		if (ln->ev_t[0])
			lz4e_b200_event_record(ctx, ln->ev_t[0], ln->stream);
	}
	/* tables first: the transfer kernel reads them */
	if (!rc)
		rc = lz4e_b200_copy_h2d(ctx, ln->stream, ln->dev, ln->pin, ln->o_ret);
	if (!rc && !(flags & LZ4E_B200_F_SRC_RESIDENT) && side_in_arenas(&ln->src) && row_w) {
		const int k = ln->src.reg_arena;

		rc = lz4e_b200_copy2d_h2d(ctx, ln->stream, as->v[k].dev + ln->src.reg_first,
					  ln->src.reg_pitch, as->v[k].host + ln->src.reg_first,
					  ln->src.reg_pitch, row_w, n);
		if (!rc && ln->src.max_size > row_w)
			rc = lz4e_b200_xfer_skip_dev(ctx, ln->stream, d_sbv, d_sit, NULL, (uint32_t)n, 0,
						     row_w);
		ws->trace.rows_up++;
	} else if (!rc && !(flags & LZ4E_B200_F_SRC_RESIDENT) && side_in_arenas(&ln->src)) {
		if (src_kernel) {
			rc = lz4e_b200_xfer_dev(ctx, ln->stream, d_sbv, d_sit, NULL, (uint32_t)n, 0);
			ws->trace.kern_up++;
		} else {
			rc = copy_ranges(ctx, as, ln->stream, ln->src.ar, 1);
			ws->trace.bulk_up++;
		}
	}
	if (!rc && !dst_kernel && (flags & LZ4E_B200_F_DST_PRESERVE))
		rc = copy_ranges(ctx, as, ln->stream, ln->dst.ar, 1);
	if (ws->tracing >= 2 && ln->ev_t[1])
		lz4e_b200_event_record(ctx, ln->ev_t[1], ln->stream);
	if (!rc && prev_kernel)
		rc = lz4e_b200_stream_wait_event(ctx, ln->stream, prev_kernel);
	if (rc)
		return rc;
	if (decompress) {
		rc = lz4e_b200_decompress_dev(
			ctx, ln->stream, d_sbv, d_sit, d_dbv, d_dit, (uint32_t)n,
			ln->src.max_size, ln->dst.max_size,
			ln->dst.all_linear ? LZ4E_B200_DEV_DST_LINEAR : 0,
			(int32_t *)(ln->dev + ln->o_ret));
	} else {
		rc = lz4e_b200_compress_dev(
			ctx, ln->stream, d_sbv, d_sit, d_dbv, d_dit, (uint32_t)n,
			ln->src.max_size,
			(ln->src.any_u16 ? 0 : LZ4E_B200_DEV_NO_BYU16) |
				(ln->dst.all_linear ? LZ4E_B200_DEV_DST_LINEAR : 0),
			(int32_t *)(ln->dev + ln->o_ret),
			(struct lz4e_b200_dev_end *)(ln->dev + ln->o_end));
	}
	if (!rc)
		rc = lz4e_b200_event_record(ctx, ln->ev_kernel, ln->stream);
	if (ws->tracing >= 2 && ln->ev_t[2])
		lz4e_b200_event_record(ctx, ln->ev_t[2], ln->stream);
	if (!rc && !(flags & LZ4E_B200_F_DST_ON_DEVICE)) {
		if (dst_kernel) {
			rc = lz4e_b200_xfer_dev(ctx, ln->stream, d_dbv, d_dit,
						(const int32_t *)(ln->dev + ln->o_ret), (uint32_t)n, 1);
			ws->trace.kern_down++;
		} else {
			rc = copy_ranges(ctx, as, ln->stream, ln->dst.ar, 0);
			ws->trace.bulk_down++;
		}
	}
	if (!rc)
		rc = lz4e_b200_copy_d2h(ctx, ln->stream, ln->pin + ln->o_ret, ln->dev + ln->o_ret,
					pin_bytes - ln->o_ret);
	if (ws->tracing >= 2 && ln->ev_t[3])
		lz4e_b200_event_record(ctx, ln->ev_t[3], ln->stream);
	ln->busy = 1; /* something may be enqueued even when rc != 0 */
	ws->trace.enqueue += now_ms() - t1;
	return rc;
}

/* bytes (source + destination capacity) per chunk: LZ4E_B200_CHUNK_MB, else a
 * sixteenth of the batch within [16, 256] MiB — small chunks pay per-chunk
 * launch/sync latency, large ones a longer pipeline fill */
static size_t chunk_target_bytes(uint64_t total)
{
	const char *e = getenv("LZ4E_B200_CHUNK_MB");
	uint64_t t = total / 16;

	if (e && atol(e) > 0)
		return (size_t)atol(e) << 20;
	if (t < ((uint64_t)16 << 20))
		t = (uint64_t)16 << 20;
	if (t > ((uint64_t)256 << 20))
		t = (uint64_t)256 << 20;
	return (size_t)t;
}

static int run_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs, size_t n,
		     unsigned flags, int decompress)
{
	struct arena_set as;
	struct host_ws *ws;
	size_t i, first, nchunk = 0, target;
	uint64_t total = 0, done_bytes = 0;
	void *prev_kernel = NULL;
	int rc = LZ4E_B200_OK, l;
	double t_start = 0;

	if (!ctx) {
		fprintf(stderr, "lz4e_b200: no CUDA context: the LZ4E library has no CPU path\n");
		for (i = 0; i < n; i++)
			reqs[i].ret = decompress ? -1 : 0;
		return LZ4E_B200_ENODEV;
	}
	if (n == 0)
		return LZ4E_B200_OK;
	if (n > 0x7fffffffu)
		return LZ4E_B200_EINVAL;
	ws = ws_acquire(ctx);
	if (!ws) {
		for (i = 0; i < n; i++)
			reqs[i].ret = decompress ? -1 : 0;
		return LZ4E_B200_EIO;
	}
	t_start = now_ms();
	memset(&ws->trace, 0, sizeof(ws->trace));
	{
		const char *e = getenv("LZ4E_B200_TRACE");

		ws->tracing = e ? atoi(e) : 0;
		ws->t_batch = t_start;
		if (ws->tracing >= 2) {
			if (!ws->ev_batch)
				ws->ev_batch = lz4e_b200_event_create_timed(ctx);
			if (ws->ev_batch)
				lz4e_b200_event_record(ctx, ws->ev_batch, ws->lane[0].stream);
			else
				ws->tracing = 1;
		}
	}
	arena_set_init(ctx, &as);
	{
		/* size of the batch, estimated from a sample: walking a quarter of
		 * a million requests before the first chunk is submitted costs a
		 * millisecond of idle PCIe */
		const size_t stride = n > 4096 ? n / 1024 : 1;
		size_t cnt = 0;

		for (i = 0; i < n; i += stride, cnt++)
			total += (uint64_t)reqs[i].src_iter->bi_size + reqs[i].dst_iter->bi_size;
		total = total / cnt * n;
	}
	target = chunk_target_bytes(total);
	/* few blocks: one kernel (each block is one warp's serial work, so two
	 * half-empty kernels one after the other only double the latency) */
	if (n < 4096 && !getenv("LZ4E_B200_CHUNK_MB"))
		target = (size_t)-1;
	if (xfer_mode() == 0 && total > target) {
		/* copy engines only: ranges of different chunks may cross; the
		 * round-1 rule — one chunk unless sources and destinations are
		 * apart — is kept for this diagnostic mode by running one chunk */
		target = (size_t)-1;
	}
	for (first = 0; first < n && !rc;) {
		struct lane *ln = &ws->lane[nchunk % NLANES];
		size_t cnt = 0, ns = 0, nd = 0;
		uint64_t bytes = 0;
		/* the first chunks are smaller so that the first kernel starts
		 * (and the last download ends) sooner */
		size_t want = nchunk == 0 && target != (size_t)-1 ? target / 4
			      : nchunk == 1 && target != (size_t)-1 ? target / 2
								    : target;

		/* ... and the last ones shrink again: what is behind the last
		 * upload (kernel, results down) is not overlapped by anything */
		if (target != (size_t)-1 && nchunk >= 2) {
			const uint64_t left = total > done_bytes ? total - done_bytes : 0;

			if (left < 2 * (uint64_t)target)
				want = left / 2 > target / 4 ? (size_t)(left / 2) : target / 4;
		}

		struct span sp = { (uintptr_t)-1, 0, 0 };
		int stage_src;

		while (first + cnt < n && (cnt == 0 || (bytes < want && cnt < (1u << 20)))) {
			const struct lz4e_b200_req *q = &reqs[first + cnt];

			ns += scan_bvecs(q->src, q->src_iter, &sp);
			nd += count_bvecs(q->dst, q->dst_iter);
			bytes += (uint64_t)q->src_iter->bi_size + q->dst_iter->bi_size;
			cnt++;
		}
		/*
		 * Sources that do not (nearly) tile the address range they span —
		 * compressed blocks in LZ4E_COMPRESSBOUND-sized slots, segments
		 * with foreign pages in between — are packed into pinned staging by
		 * the host threads and go up in one copy: exactly their bytes, at
		 * copy-engine speed.  (SM reads over PCIe, the transfer kernel in
		 * the host -> device direction, stop at about 30 GB/s and stall
		 * behind concurrent device -> host writes; it stays as the path for
		 * sides that turn out sparse after a dense guess.)
		 */
		stage_src = xfer_mode() == 1 && sparse_src_mode(NULL) == SPARSE_PACK &&
			    !(flags & LZ4E_B200_F_SRC_RESIDENT) && sp.sum &&
			    (double)sp.sum < 0.97 * (double)(sp.hi - sp.lo);
		rc = lane_finish(ctx, ws, ln, decompress, flags);
		ln->chunk_no = nchunk;
		ln->chunk_bytes = (size_t)bytes;
		if (!rc)
			rc = lane_submit(ctx, ws, &as, ln, reqs + first, cnt, ns, nd, flags,
					 decompress, prev_kernel, stage_src);
		prev_kernel = ln->ev_kernel;
		done_bytes += bytes;
		first += cnt;
		nchunk++;
	}
	/* drain in submission order */
	for (l = 0; l < NLANES; l++) {
		struct lane *ln = &ws->lane[(nchunk + l) % NLANES];
		const int r2 = lane_finish(ctx, ws, ln, decompress, flags);

		if (!rc)
			rc = r2;
	}
	if (ws->tracing)
		fprintf(stderr,
			"lz4e_b200: dev %d %s batch of %zu in %zu chunks: %.2f ms (table build %.2f, "
			"packing %.2f, enqueue %.2f, waiting %.2f); source up: %zu chunks "
			"by copy engine, %zu as rows + excess, %zu packed + copy engine, %zu by kernel; results down: %zu "
			"by copy engine, %zu by kernel\n",
			lz4e_b200_ctx_device(ctx), decompress ? "decompress" : "compress", n, nchunk,
			now_ms() - t_start, ws->trace.build, ws->trace.pack,
			ws->trace.enqueue - ws->trace.pack,
			ws->trace.wait, ws->trace.bulk_up, ws->trace.rows_up, ws->trace.packed_up,
			ws->trace.kern_up,
			ws->trace.bulk_down, ws->trace.kern_down);
	if (rc) {
		fprintf(stderr, "lz4e_b200: batch failed (%d): %s\n", rc,
			lz4e_b200_strerror(ctx));
		for (i = 0; i < n; i++)
			reqs[i].ret = decompress ? -1 : 0;
	}
	pthread_mutex_unlock(&ws->mu);
	return rc;
}

int lz4e_b200_compress_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs,
			     size_t n, unsigned flags)
{
	return run_batch(ctx, reqs, n, flags, 0);
}

int lz4e_b200_decompress_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs,
			       size_t n, unsigned flags)
{
	return run_batch(ctx, reqs, n, flags | LZ4E_B200_F_KEEP_ITERS, 1);
}

/* ---- the reference's own entry points ------------------------------------ */

int LZ4E_compress_default(const struct bio_vec *src, struct bio_vec *dst,
			  struct bvec_iter *srcIter, struct bvec_iter *dstIter,
			  void *wrkmem)
{
	struct lz4e_b200_req r;

	if (wrkmem) /* the reference clears its state first, lz4e_compress.c:548 */
		memset(wrkmem, 0, LZ4E_MEM_COMPRESS);
	r.src = src;
	r.dst = dst;
	r.src_iter = srcIter;
	r.dst_iter = dstIter;
	r.ret = 0;
	r.rsvd = 0;
	/* a lone call must not replace bytes of destination pages it does not
	 * write, like the reference */
	if (run_batch(lz4e_b200_default_ctx(), &r, 1, LZ4E_B200_F_DST_PRESERVE, 0))
		return 0;
	return r.ret;
}

int LZ4E_decompress_safe_sg(const struct bio_vec *src, struct bio_vec *dst,
			    const struct bvec_iter *srcIter,
			    const struct bvec_iter *dstIter)
{
	struct lz4e_b200_req r;
	struct bvec_iter si = *srcIter, di = *dstIter;

	r.src = src;
	r.dst = dst;
	r.src_iter = &si;
	r.dst_iter = &di;
	r.ret = -1;
	r.rsvd = 0;
	if (run_batch(lz4e_b200_default_ctx(), &r, 1,
		      LZ4E_B200_F_DST_PRESERVE | LZ4E_B200_F_KEEP_ITERS, 1))
		return -1;
	return r.ret;
}

/* describe a contiguous host buffer as one (multi-page) bio_vec */
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

int LZ4E_decompress_safe(const char *source, char *dest, int compressedSize,
			 int maxDecompressedSize)
{
	struct bio_vec sbv, dbv;
	struct bvec_iter si, di;

	if (compressedSize < 0 || maxDecompressedSize < 0)
		return -1;
	buf_to_bvec(source, (unsigned int)compressedSize, &sbv, &si);
	buf_to_bvec(dest, (unsigned int)maxDecompressedSize, &dbv, &di);
	return LZ4E_decompress_safe_sg(&sbv, &dbv, &si, &di);
}
