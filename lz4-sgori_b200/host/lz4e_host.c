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

#include "../../include/lz4e.h"
#include "../../include/lz4e_b200.h"

#define MAX_ARENAS 16

struct arena_range {
	lz4e_b200_arena *a;
	size_t lo, hi; /* byte range touched inside the arena */
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
	struct arena_range ar[MAX_ARENAS];
	int nar;
	struct bounce *bn;
	size_t nbn, cap_bn;
	size_t bounce_bytes;
	uint32_t max_size;
	int all_linear; /* every block is a single segment */
	int any_u16;	/* some block would select byU16 (compress source) */
};

/* one workspace per process, serialised: a lazily created resource pool
 * behind a lock (SURVEY.md section 8b "Threading") */
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static struct {
	lz4e_b200_ctx *ctx;
	uint8_t *pin, *dev, *bounce;
	size_t pin_bytes, dev_bytes, bounce_dev_bytes;
} g_ws;

static int ws_reserve(lz4e_b200_ctx *ctx, size_t pin_bytes, size_t bounce_bytes)
{
	if (g_ws.ctx != ctx) {
		if (g_ws.ctx) {
			lz4e_b200_host_free(g_ws.ctx, g_ws.pin);
			lz4e_b200_dev_free(g_ws.ctx, g_ws.dev);
			lz4e_b200_dev_free(g_ws.ctx, g_ws.bounce);
		}
		memset(&g_ws, 0, sizeof(g_ws));
		g_ws.ctx = ctx;
	}
	if (pin_bytes > g_ws.pin_bytes) {
		size_t want = pin_bytes + pin_bytes / 4 + 4096;

		lz4e_b200_sync(ctx, NULL);
		lz4e_b200_host_free(ctx, g_ws.pin);
		lz4e_b200_dev_free(ctx, g_ws.dev);
		g_ws.pin = lz4e_b200_host_alloc(ctx, want);
		g_ws.dev = lz4e_b200_dev_alloc(ctx, want);
		g_ws.pin_bytes = g_ws.dev_bytes = (g_ws.pin && g_ws.dev) ? want : 0;
		if (!g_ws.pin_bytes)
			return LZ4E_B200_ENOMEM;
	}
	if (bounce_bytes > g_ws.bounce_dev_bytes) {
		size_t want = bounce_bytes + bounce_bytes / 4 + 4096;

		lz4e_b200_sync(ctx, NULL);
		lz4e_b200_dev_free(ctx, g_ws.bounce);
		g_ws.bounce = lz4e_b200_dev_alloc(ctx, want);
		g_ws.bounce_dev_bytes = g_ws.bounce ? want : 0;
		if (!g_ws.bounce)
			return LZ4E_B200_ENOMEM;
	}
	return LZ4E_B200_OK;
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

static void range_add(struct side *s, lz4e_b200_arena *a, size_t lo, size_t hi)
{
	int i;

	for (i = 0; i < s->nar; i++) {
		if (s->ar[i].a == a) {
			if (lo < s->ar[i].lo)
				s->ar[i].lo = lo;
			if (hi > s->ar[i].hi)
				s->ar[i].hi = hi;
			return;
		}
	}
	if (s->nar < MAX_ARENAS) {
		s->ar[s->nar].a = a;
		s->ar[s->nar].lo = lo;
		s->ar[s->nar].hi = hi;
		s->nar++;
	}
}

/* Append the bio_vecs block `blk` covers to side s.  The device iterator is
 * rebased so that its bi_idx is 0 (indices relative to the start are all the
 * reference's table-type rule looks at, lz4e_compress.c:195-202). */
static int side_add(lz4e_b200_ctx *ctx, struct side *s, uint32_t blk,
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
		lz4e_b200_arena *a = NULL;
		uint64_t d;
		struct lz4e_b200_dev_bvec *o = &s->bv[s->nbv++];

		if (take > left)
			take = left;
		if (nseg >= 16 || v->bv_len > 4096)
			u16 = 0;
		d = lz4e_b200_arena_translate(ctx, h + done, &a);
		if (d && lz4e_b200_arena_translate(ctx, h + done + take - 1, NULL)) {
			const size_t off = (size_t)((h + done) -
						    (uint8_t *)lz4e_b200_arena_pages(a));

			o->addr = d - done; /* address of the bvec's byte 0 */
			o->rsvd = 0;
			range_add(s, a, off, off + take);
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

static size_t al256(size_t x)
{
	return (x + 255) & ~(size_t)255;
}

static int run_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs, size_t n,
		     unsigned flags, int decompress)
{
	struct side src, dst;
	size_t i, ns = 0, nd = 0, off, pin_bytes;
	size_t o_sbv, o_dbv, o_sit, o_dit, o_ret, o_end;
	int32_t *rets;
	struct lz4e_b200_dev_end *ends;
	int rc = LZ4E_B200_OK;

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

	memset(&src, 0, sizeof(src));
	memset(&dst, 0, sizeof(dst));
	src.all_linear = dst.all_linear = 1;
	for (i = 0; i < n; i++) {
		ns += count_bvecs(reqs[i].src, reqs[i].src_iter);
		nd += count_bvecs(reqs[i].dst, reqs[i].dst_iter);
	}
	off = 0;
	o_sbv = off; off = al256(off + ns * sizeof(struct lz4e_b200_dev_bvec));
	o_dbv = off; off = al256(off + nd * sizeof(struct lz4e_b200_dev_bvec));
	o_sit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	o_dit = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_iter));
	o_ret = off; off = al256(off + n * sizeof(int32_t));
	o_end = off; off = al256(off + n * sizeof(struct lz4e_b200_dev_end));
	pin_bytes = off;

	pthread_mutex_lock(&g_mu);
	rc = ws_reserve(ctx, pin_bytes, 0);
	if (rc)
		goto out;
	src.bv = (struct lz4e_b200_dev_bvec *)(g_ws.pin + o_sbv);
	dst.bv = (struct lz4e_b200_dev_bvec *)(g_ws.pin + o_dbv);
	src.it = (struct lz4e_b200_dev_iter *)(g_ws.pin + o_sit);
	dst.it = (struct lz4e_b200_dev_iter *)(g_ws.pin + o_dit);
	rets = (int32_t *)(g_ws.pin + o_ret);
	ends = (struct lz4e_b200_dev_end *)(g_ws.pin + o_end);
	for (i = 0; i < n && !rc; i++) {
		rc = side_add(ctx, &src, (uint32_t)i, reqs[i].src, reqs[i].src_iter);
		if (!rc)
			rc = side_add(ctx, &dst, (uint32_t)i, reqs[i].dst, reqs[i].dst_iter);
	}
	if (rc)
		goto out;
	/* bounce space: source segments first, destination segments after */
	{
		const size_t sb = (src.bounce_bytes + 255) & ~(size_t)255;

		rc = ws_reserve(ctx, pin_bytes, sb + dst.bounce_bytes + 256);
		if (rc)
			goto out;
		side_rebase_bounce(&src, g_ws.bounce);
		side_rebase_bounce(&dst, g_ws.bounce + sb);
		for (i = 0; i < src.nbn && !rc; i++)
			rc = lz4e_b200_copy_h2d(ctx, NULL, g_ws.bounce + src.bn[i].dev_off,
						src.bn[i].host, src.bn[i].len);
		for (i = 0; i < dst.nbn; i++)
			dst.bn[i].dev_off += sb;
	}
	/* pages of arenas: one bulk copy per arena and direction */
	if (!(flags & LZ4E_B200_F_SRC_RESIDENT)) {
		int k;

		for (k = 0; k < src.nar && !rc; k++) {
			uint8_t *hb = (uint8_t *)lz4e_b200_arena_pages(src.ar[k].a);
			uint8_t *db = (uint8_t *)lz4e_b200_arena_dev_base(src.ar[k].a);

			rc = lz4e_b200_copy_h2d(ctx, NULL, db + src.ar[k].lo,
						hb + src.ar[k].lo,
						src.ar[k].hi - src.ar[k].lo);
		}
	}
	if (flags & LZ4E_B200_F_DST_PRESERVE) {
		int k;

		for (k = 0; k < dst.nar && !rc; k++) {
			uint8_t *hb = (uint8_t *)lz4e_b200_arena_pages(dst.ar[k].a);
			uint8_t *db = (uint8_t *)lz4e_b200_arena_dev_base(dst.ar[k].a);

			rc = lz4e_b200_copy_h2d(ctx, NULL, db + dst.ar[k].lo,
						hb + dst.ar[k].lo,
						dst.ar[k].hi - dst.ar[k].lo);
		}
	}
	if (!rc)
		rc = lz4e_b200_copy_h2d(ctx, NULL, g_ws.dev, g_ws.pin, o_ret);
	if (rc)
		goto out;
	if (decompress) {
		rc = lz4e_b200_decompress_dev(
			ctx, NULL, (struct lz4e_b200_dev_bvec *)(g_ws.dev + o_sbv),
			(struct lz4e_b200_dev_iter *)(g_ws.dev + o_sit),
			(struct lz4e_b200_dev_bvec *)(g_ws.dev + o_dbv),
			(struct lz4e_b200_dev_iter *)(g_ws.dev + o_dit), (uint32_t)n,
			src.max_size, dst.max_size,
			dst.all_linear ? LZ4E_B200_DEV_DST_LINEAR : 0,
			(int32_t *)(g_ws.dev + o_ret));
	} else {
		rc = lz4e_b200_compress_dev(
			ctx, NULL, (struct lz4e_b200_dev_bvec *)(g_ws.dev + o_sbv),
			(struct lz4e_b200_dev_iter *)(g_ws.dev + o_sit),
			(struct lz4e_b200_dev_bvec *)(g_ws.dev + o_dbv),
			(struct lz4e_b200_dev_iter *)(g_ws.dev + o_dit), (uint32_t)n,
			src.max_size, src.any_u16 ? 0 : LZ4E_B200_DEV_NO_BYU16,
			(int32_t *)(g_ws.dev + o_ret),
			(struct lz4e_b200_dev_end *)(g_ws.dev + o_end));
	}
	if (rc)
		goto out;
	rc = lz4e_b200_copy_d2h(ctx, NULL, g_ws.pin + o_ret, g_ws.dev + o_ret,
				pin_bytes - o_ret);
	if (!rc && !(flags & LZ4E_B200_F_DST_ON_DEVICE)) {
		int k;

		for (k = 0; k < dst.nar && !rc; k++) {
			uint8_t *hb = (uint8_t *)lz4e_b200_arena_pages(dst.ar[k].a);
			uint8_t *db = (uint8_t *)lz4e_b200_arena_dev_base(dst.ar[k].a);

			rc = lz4e_b200_copy_d2h(ctx, NULL, hb + dst.ar[k].lo,
						db + dst.ar[k].lo,
						dst.ar[k].hi - dst.ar[k].lo);
		}
	}
	if (!rc)
		rc = lz4e_b200_sync(ctx, NULL);
	if (rc)
		goto out;
	/* bounced destination segments: bring back exactly the bytes produced */
	for (i = 0; i < dst.nbn && !rc; i++) {
		const struct bounce *b = &dst.bn[i];
		const int32_t r = rets[b->blk];
		uint32_t len;

		if (r <= 0 || (uint32_t)r <= b->pos)
			continue;
		len = (uint32_t)r - b->pos;
		if (len > b->len)
			len = b->len;
		rc = lz4e_b200_copy_d2h(ctx, NULL, b->host, g_ws.bounce + b->dev_off, len);
	}
	if (!rc && dst.nbn)
		rc = lz4e_b200_sync(ctx, NULL);
	if (rc)
		goto out;
	for (i = 0; i < n; i++) {
		reqs[i].ret = rets[i];
		if (!decompress && rets[i] > 0 && !(flags & LZ4E_B200_F_KEEP_ITERS)) {
			/* leave the iterators where the reference does,
			 * lz4e_compress.c:304-305 and :528-529 */
			lz4e_bvec_iter_advance(reqs[i].src, reqs[i].src_iter,
					       ends[i].src_pos);
			lz4e_bvec_iter_advance(reqs[i].dst, reqs[i].dst_iter,
					       ends[i].dst_pos);
		}
	}
out:
	if (rc) {
		fprintf(stderr, "lz4e_b200: batch failed (%d): %s\n", rc,
			lz4e_b200_strerror(ctx));
		for (i = 0; i < n; i++)
			reqs[i].ret = decompress ? -1 : 0;
	}
	pthread_mutex_unlock(&g_mu);
	free(src.bn);
	free(dst.bn);
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
