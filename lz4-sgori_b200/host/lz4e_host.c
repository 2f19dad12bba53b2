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
	struct bounce *bn;
	size_t nbn, cap_bn;
	size_t bounce_bytes;
	uint32_t max_size;
	int all_linear; /* every block is a single segment */
	int any_u16;	/* some block would select byU16 (compress source) */
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
	uint8_t *pin, *dev, *bounce;
	size_t pin_bytes, bounce_bytes;
	int busy;
	struct lz4e_b200_req *reqs;
	size_t n;
	struct side src, dst;
	size_t o_ret, o_end;
};

/* one workspace per process, serialised: a lazily created resource pool
 * behind a lock (SURVEY.md section 8b "Threading") */
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static struct {
	lz4e_b200_ctx *ctx;
	struct lane lane[NLANES];
} g_ws;

static void ws_drop(void)
{
	int l;

	if (!g_ws.ctx)
		return;
	for (l = 0; l < NLANES; l++) {
		struct lane *ln = &g_ws.lane[l];

		lz4e_b200_host_free(g_ws.ctx, ln->pin);
		lz4e_b200_dev_free(g_ws.ctx, ln->dev);
		lz4e_b200_dev_free(g_ws.ctx, ln->bounce);
		if (ln->ev_kernel)
			lz4e_b200_event_destroy(g_ws.ctx, ln->ev_kernel);
		if (ln->stream)
			lz4e_b200_stream_destroy(g_ws.ctx, ln->stream);
		free(ln->src.bn);
		free(ln->dst.bn);
	}
	memset(&g_ws, 0, sizeof(g_ws));
}

static int ws_bind(lz4e_b200_ctx *ctx)
{
	int l;

	if (g_ws.ctx == ctx)
		return LZ4E_B200_OK;
	ws_drop();
	g_ws.ctx = ctx;
	for (l = 0; l < NLANES; l++) {
		g_ws.lane[l].stream = lz4e_b200_stream_create(ctx);
		g_ws.lane[l].ev_kernel = lz4e_b200_event_create(ctx);
		if (!g_ws.lane[l].stream || !g_ws.lane[l].ev_kernel)
			return LZ4E_B200_EIO;
	}
	return LZ4E_B200_OK;
}

/* the lane is idle when this is called */
static int lane_reserve(lz4e_b200_ctx *ctx, struct lane *ln, size_t pin_bytes,
			size_t bounce_bytes)
{
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

/* arena byte ranges a block's segments touch (no tables built) */
static void side_ranges(struct arena_set *as, struct range *r,
			const struct bio_vec *bv, const struct bvec_iter *it)
{
	unsigned int left = it->bi_size, idx = it->bi_idx, done = it->bi_bvec_done;

	while (left) {
		const struct bio_vec *v = &bv[idx];
		unsigned int take = v->bv_len - done;
		uint8_t *h = (uint8_t *)page_address(v->bv_page) + v->bv_offset + done;
		int k;

		if (take > left)
			take = left;
		k = arena_find(as, h, take);
		if (k >= 0)
			range_add(&r[k], (size_t)(h - as->v[k].host),
				  (size_t)(h - as->v[k].host) + take);
		left -= take;
		done = 0;
		idx++;
	}
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
		k = arena_find(as, h + done, take);
		if (k >= 0) {
			const size_t off = (size_t)((h + done) - as->v[k].host);

			/* address of the bvec's byte 0 */
			o->addr = (uint64_t)(uintptr_t)(as->v[k].dev + off) - done;
			o->rsvd = 0;
			range_add(&s->ar[k], off, off + take);
		} else {
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
	if (nseg > 1)
		s->all_linear = 0;
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

/* LZ4E_B200_TRACE=1: where a batch spends its host time (stderr) */
static struct {
	double build, enqueue, wait;
} g_trace;

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

/* wait for a chunk and hand its results back to the caller's requests */
static int lane_finish(lz4e_b200_ctx *ctx, struct lane *ln, int decompress,
		       unsigned flags)
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
	g_trace.wait += now_ms() - t0;
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
		if (!decompress && rets[i] > 0 && !(flags & LZ4E_B200_F_KEEP_ITERS)) {
			/* leave the iterators where the reference does,
			 * lz4e_compress.c:304-305 and :528-529 */
			lz4e_bvec_iter_advance(q->src, q->src_iter, ends[i].src_pos);
			lz4e_bvec_iter_advance(q->dst, q->dst_iter, ends[i].dst_pos);
		}
	}
	return LZ4E_B200_OK;
}

/* build the tables of one chunk in the lane's workspace and enqueue its
 * uploads, kernel and downloads on the lane's stream */
static int lane_submit(lz4e_b200_ctx *ctx, struct arena_set *as, struct lane *ln,
		       struct lz4e_b200_req *reqs, size_t n, size_t ns, size_t nd,
		       unsigned flags, int decompress, void *prev_kernel)
{
	size_t i, off, pin_bytes, o_sbv, o_dbv, o_sit, o_dit, sb;
	int rc = LZ4E_B200_OK, l;
	double t0 = now_ms(), t1;

	off = 0;
	o_sbv = off; off = al256(off + ns * sizeof(struct lz4e_b200_dev_bvec));
	o_dbv = off; off = al256(off + nd * sizeof(struct lz4e_b200_dev_bvec));
	o_sit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	o_dit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	ln->o_ret = off; off = al256(off + n * sizeof(int32_t));
	ln->o_end = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_end));
	pin_bytes = off;
	rc = lane_reserve(ctx, ln, pin_bytes, 0);
	if (rc)
		return rc;
	ln->reqs = reqs;
	ln->n = n;
	side_reset(&ln->src);
	side_reset(&ln->dst);
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
	t1 = now_ms();
	g_trace.build += t1 - t0;
	/*
	 * Chunks still in flight: a bulk copy of this chunk must not cross bytes a
	 * running chunk produces (or the other way round).  Callers that interleave
	 * the pages of different chunks inside one arena lose the overlap, not the
	 * result.
	 */
	for (l = 0; l < NLANES && !rc; l++) {
		struct lane *o = &g_ws.lane[l];

		if (o == ln || !o->busy)
			continue;
		if (ranges_overlap(ln->src.ar, o->dst.ar) || ranges_overlap(ln->dst.ar, o->dst.ar) ||
		    ranges_overlap(ln->dst.ar, o->src.ar))
			rc = lane_finish(ctx, o, decompress, flags);
	}
	if (rc)
		return rc;
	/* bounce space: source segments first, destination segments after */
	sb = (ln->src.bounce_bytes + 255) & ~(size_t)255;
	if (ln->src.nbn || ln->dst.nbn) {
		rc = lane_reserve(ctx, ln, pin_bytes, sb + ln->dst.bounce_bytes + 256);
		if (rc)
			return rc;
		side_rebase_bounce(&ln->src, ln->bounce);
		side_rebase_bounce(&ln->dst, ln->bounce + sb);
		for (i = 0; i < ln->src.nbn && !rc; i++)
			rc = lz4e_b200_copy_h2d(ctx, ln->stream, ln->bounce + ln->src.bn[i].dev_off,
						ln->src.bn[i].host, ln->src.bn[i].len);
		for (i = 0; i < ln->dst.nbn; i++)
			ln->dst.bn[i].dev_off += sb;
	}
	/* pages of arenas: one bulk copy per arena and direction */
	if (!rc && !(flags & LZ4E_B200_F_SRC_RESIDENT))
		rc = copy_ranges(ctx, as, ln->stream, ln->src.ar, 1);
	if (!rc && (flags & LZ4E_B200_F_DST_PRESERVE))
		rc = copy_ranges(ctx, as, ln->stream, ln->dst.ar, 1);
	if (!rc)
		rc = lz4e_b200_copy_h2d(ctx, ln->stream, ln->dev, ln->pin, ln->o_ret);
	if (!rc && prev_kernel)
		rc = lz4e_b200_stream_wait_event(ctx, ln->stream, prev_kernel);
	if (rc)
		return rc;
	if (decompress) {
		rc = lz4e_b200_decompress_dev(
			ctx, ln->stream, (struct lz4e_b200_dev_bvec *)(ln->dev + o_sbv),
			(struct lz4e_b200_dev_iter *)(ln->dev + o_sit),
			(struct lz4e_b200_dev_bvec *)(ln->dev + o_dbv),
			(struct lz4e_b200_dev_iter *)(ln->dev + o_dit), (uint32_t)n,
			ln->src.max_size, ln->dst.max_size,
			ln->dst.all_linear ? LZ4E_B200_DEV_DST_LINEAR : 0,
			(int32_t *)(ln->dev + ln->o_ret));
	} else {
		rc = lz4e_b200_compress_dev(
			ctx, ln->stream, (struct lz4e_b200_dev_bvec *)(ln->dev + o_sbv),
			(struct lz4e_b200_dev_iter *)(ln->dev + o_sit),
			(struct lz4e_b200_dev_bvec *)(ln->dev + o_dbv),
			(struct lz4e_b200_dev_iter *)(ln->dev + o_dit), (uint32_t)n,
			ln->src.max_size,
			(ln->src.any_u16 ? 0 : LZ4E_B200_DEV_NO_BYU16) |
				(ln->dst.all_linear ? LZ4E_B200_DEV_DST_LINEAR : 0),
			(int32_t *)(ln->dev + ln->o_ret),
			(struct lz4e_b200_dev_end *)(ln->dev + ln->o_end));
	}
	if (!rc)
		rc = lz4e_b200_event_record(ctx, ln->ev_kernel, ln->stream);
	if (!rc)
		rc = lz4e_b200_copy_d2h(ctx, ln->stream, ln->pin + ln->o_ret, ln->dev + ln->o_ret,
					pin_bytes - ln->o_ret);
	if (!rc && !(flags & LZ4E_B200_F_DST_ON_DEVICE))
		rc = copy_ranges(ctx, as, ln->stream, ln->dst.ar, 0);
	ln->busy = 1; /* something may be enqueued even when rc != 0 */
	g_trace.enqueue += now_ms() - t1;
	return rc;
}

/* bytes (source + destination capacity) per chunk: LZ4E_B200_CHUNK_MB, else a
 * sixteenth of the batch within [16, 256] MiB — small chunks pay per-chunk
 * launch/sync latency, large ones a longer pipeline fill (B200, PCIe 5: 1 GiB of
 * 4 KiB blocks takes 38 ms at 32 MiB, 33 ms at 128 MiB) */
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
	size_t i, first, nchunk = 0, target;
	uint64_t total = 0;
	void *prev_kernel = NULL;
	int rc = LZ4E_B200_OK, l;
	double t_start = 0, t_guard = 0;

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

	pthread_mutex_lock(&g_mu);
	t_start = now_ms();
	memset(&g_trace, 0, sizeof(g_trace));
	rc = ws_bind(ctx);
	if (rc)
		goto out;
	arena_set_init(ctx, &as);
	for (i = 0; i < n; i++)
		total += (uint64_t)reqs[i].src_iter->bi_size + reqs[i].dst_iter->bi_size;
	target = chunk_target_bytes(total);
	/* few blocks: one kernel (each block is one warp's serial work, so two
	 * half-empty kernels one after the other only double the latency) */
	if (n < 4096 && !getenv("LZ4E_B200_CHUNK_MB"))
		target = (size_t)-1;
	if (total > target) {
		/*
		 * More than one chunk: uploads of later chunks and downloads of
		 * earlier ones will overlap in time.  That is only safe when no bulk
		 * source range crosses a bulk destination range of the same arena;
		 * otherwise the batch runs as a single chunk (all uploads, kernel,
		 * all downloads).
		 */
		struct range sr[MAX_ARENAS], dr[MAX_ARENAS];

		memset(sr, 0, sizeof(sr));
		memset(dr, 0, sizeof(dr));
		for (i = 0; i < n; i++) {
			side_ranges(&as, sr, reqs[i].src, reqs[i].src_iter);
			side_ranges(&as, dr, reqs[i].dst, reqs[i].dst_iter);
		}
		if (ranges_overlap(sr, dr))
			target = (size_t)-1;
	}
	t_guard = now_ms() - t_start;
	for (first = 0; first < n && !rc;) {
		struct lane *ln = &g_ws.lane[nchunk % NLANES];
		size_t cnt = 0, ns = 0, nd = 0;
		uint64_t bytes = 0;

		while (first + cnt < n && (cnt == 0 || (bytes < target && cnt < (1u << 20)))) {
			const struct lz4e_b200_req *q = &reqs[first + cnt];

			ns += count_bvecs(q->src, q->src_iter);
			nd += count_bvecs(q->dst, q->dst_iter);
			bytes += (uint64_t)q->src_iter->bi_size + q->dst_iter->bi_size;
			cnt++;
		}
		rc = lane_finish(ctx, ln, decompress, flags);
		if (!rc)
			rc = lane_submit(ctx, &as, ln, reqs + first, cnt, ns, nd, flags,
					 decompress, prev_kernel);
		prev_kernel = ln->ev_kernel;
		first += cnt;
		nchunk++;
	}
	/* drain in submission order */
	for (l = 0; l < NLANES; l++) {
		struct lane *ln = &g_ws.lane[(nchunk + l) % NLANES];
		const int r2 = lane_finish(ctx, ln, decompress, flags);

		if (!rc)
			rc = r2;
	}
	if (getenv("LZ4E_B200_TRACE"))
		fprintf(stderr,
			"lz4e_b200: %s batch of %zu in %zu chunks: %.2f ms (range check %.2f, "
			"table build %.2f, enqueue %.2f, waiting %.2f)\n",
			decompress ? "decompress" : "compress", n, nchunk, now_ms() - t_start,
			t_guard, g_trace.build, g_trace.enqueue, g_trace.wait);
out:
	if (rc) {
		fprintf(stderr, "lz4e_b200: batch failed (%d): %s\n", rc,
			lz4e_b200_strerror(ctx));
		for (i = 0; i < n; i++)
			reqs[i].ret = decompress ? -1 : 0;
	}
	pthread_mutex_unlock(&g_mu);
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
