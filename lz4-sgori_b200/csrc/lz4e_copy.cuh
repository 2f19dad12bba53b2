/*
 * lz4e_copy.cuh — cooperative byte copies between arbitrary byte addresses.
 *
 * These replace the reference's scatter-gather access chain
 * (LZ4E_memcpy_from_sg / LZ4E_memcpy_to_sg / LZ4E_wildCopy / LZ4E_memcpy,
 * lz4e/include/lz4e_defs.h:366-402, :547-585): instead of building a bio_vec
 * and calling memcpy per 1/2/4/8-byte access, a whole segment is moved by W
 * cooperating lanes with the widest accesses the mutual alignment allows:
 *   - source and destination congruent mod 16: 16-byte vectors,
 *   - otherwise destination-aligned 4-byte words assembled from two aligned
 *     source words with a funnel shift,
 *   - byte accesses only for the unaligned head and tail.
 * `l` is the lane's index inside its group of W lanes.  No synchronisation is
 * done here; callers order producer/consumer phases with lz4e_syncwarp().
 */
#ifndef LZ4E_COPY_CUH
#define LZ4E_COPY_CUH

#include "lz4e_simt.h"

struct lz4e_u128 {
	uint32_t x, y, z, w;
};

#if defined(__CUDACC__)
LZ4E_DEV lz4e_u128 lz4e_ld128(const uint8_t *p)
{
	uint4 v = *reinterpret_cast<const uint4 *>(p);
	lz4e_u128 r = { v.x, v.y, v.z, v.w };
	return r;
}
LZ4E_DEV void lz4e_st128(uint8_t *p, lz4e_u128 v)
{
	*reinterpret_cast<uint4 *>(p) = make_uint4(v.x, v.y, v.z, v.w);
}
#else
LZ4E_DEV lz4e_u128 lz4e_ld128(const uint8_t *p)
{
	lz4e_u128 r;
	__builtin_memcpy(&r, p, 16);
	return r;
}
LZ4E_DEV void lz4e_st128(uint8_t *p, lz4e_u128 v)
{
	__builtin_memcpy(p, &v, 16);
}
#endif

LZ4E_DEV uint32_t lz4e_ld32a(const uint8_t *p) /* p is 4-byte aligned */
{
	return *reinterpret_cast<const uint32_t *>(p);
}
LZ4E_DEV void lz4e_st32a(uint8_t *p, uint32_t v)
{
	*reinterpret_cast<uint32_t *>(p) = v;
}

/* unaligned little-endian 32-bit load built from two aligned words */
LZ4E_DEV uint32_t lz4e_ld32u(const uint8_t *p)
{
	const uintptr_t a = (uintptr_t)p;
	const uint8_t *al = (const uint8_t *)(a & ~(uintptr_t)3);
	const uint32_t sh = (uint32_t)(a & 3) * 8;

	return lz4e_funnel_r(lz4e_ld32a(al), lz4e_ld32a(al + 4), sh);
}

/*
 * Copy len bytes src -> dst with W cooperating lanes (lane index l).
 * Regions must not overlap.
 */
template <int W>
LZ4E_DEV void lz4e_copy(uint8_t *dst, const uint8_t *src, uint32_t len,
			uint32_t l)
{
	if (len < 16) {
		for (uint32_t i = l; i < len; i += W)
			dst[i] = src[i];
		return;
	}
	if ((((uintptr_t)dst ^ (uintptr_t)src) & 15) == 0) {
		const uint32_t head = (uint32_t)(-(uintptr_t)dst) & 15;
		const uint32_t body = (len - head) >> 4;
		const uint32_t tail0 = head + (body << 4);

		for (uint32_t i = l; i < head; i += W)
			dst[i] = src[i];
		uint32_t i = l;
		for (; i + 3 * W < body; i += 4 * W) {
			lz4e_u128 a = lz4e_ld128(src + head + ((size_t)i << 4));
			lz4e_u128 b = lz4e_ld128(src + head + ((size_t)(i + W) << 4));
			lz4e_u128 c = lz4e_ld128(src + head + ((size_t)(i + 2 * W) << 4));
			lz4e_u128 d = lz4e_ld128(src + head + ((size_t)(i + 3 * W) << 4));
			lz4e_st128(dst + head + ((size_t)i << 4), a);
			lz4e_st128(dst + head + ((size_t)(i + W) << 4), b);
			lz4e_st128(dst + head + ((size_t)(i + 2 * W) << 4), c);
			lz4e_st128(dst + head + ((size_t)(i + 3 * W) << 4), d);
		}
		for (; i < body; i += W)
			lz4e_st128(dst + head + ((size_t)i << 4),
				   lz4e_ld128(src + head + ((size_t)i << 4)));
		for (uint32_t t = tail0 + l; t < len; t += W)
			dst[t] = src[t];
		return;
	}
	{
		const uint32_t head = (uint32_t)(-(uintptr_t)dst) & 3;
		const uint32_t body = (len - head) >> 2;
		const uint32_t tail0 = head + (body << 2);
		const uintptr_t sa = (uintptr_t)(src + head);
		const uint8_t *al = (const uint8_t *)(sa & ~(uintptr_t)3);
		const uint32_t sh = (uint32_t)(sa & 3) * 8;

		for (uint32_t i = l; i < head; i += W) /* head <= 3 */
			dst[i] = src[i];
		if (sh == 0) {
			for (uint32_t i = l; i < body; i += W)
				lz4e_st32a(dst + head + ((size_t)i << 2),
					   lz4e_ld32a(al + ((size_t)i << 2)));
		} else {
			uint32_t i = l;
			for (; i + W < body; i += 2 * W) {
				uint32_t a0 = lz4e_ld32a(al + ((size_t)i << 2));
				uint32_t a1 = lz4e_ld32a(al + ((size_t)i << 2) + 4);
				uint32_t b0 = lz4e_ld32a(al + ((size_t)(i + W) << 2));
				uint32_t b1 = lz4e_ld32a(al + ((size_t)(i + W) << 2) + 4);
				lz4e_st32a(dst + head + ((size_t)i << 2),
					   lz4e_funnel_r(a0, a1, sh));
				lz4e_st32a(dst + head + ((size_t)(i + W) << 2),
					   lz4e_funnel_r(b0, b1, sh));
			}
			for (; i < body; i += W) {
				uint32_t a0 = lz4e_ld32a(al + ((size_t)i << 2));
				uint32_t a1 = lz4e_ld32a(al + ((size_t)i << 2) + 4);
				lz4e_st32a(dst + head + ((size_t)i << 2),
					   lz4e_funnel_r(a0, a1, sh));
			}
		}
		for (uint32_t t = tail0 + l; t < len; t += W)
			dst[t] = src[t];
	}
}

/* SG cursor: a forward-only position inside a device bvec table. */
struct lz4e_cursor {
	const struct lz4e_b200_dev_bvec *v; /* current bvec */
	uint32_t done;			    /* bytes consumed in *v */
};

LZ4E_DEV lz4e_cursor lz4e_cursor_init(const struct lz4e_b200_dev_bvec *table,
				      const struct lz4e_b200_dev_iter &it)
{
	lz4e_cursor c;

	c.v = table + it.first + it.bi_idx;
	c.done = it.bi_bvec_done;
	return c;
}

/* byte address of the cursor position; valid only while bytes remain */
LZ4E_DEV uint8_t *lz4e_cursor_addr(const lz4e_cursor &c)
{
	uint8_t *p = (uint8_t *)(uintptr_t)(c.v->addr + c.done);

	LZ4E_ASSUME_GLOBAL(p);
	return p;
}

/* move the cursor n bytes forward; the bytes must exist */
LZ4E_DEV void lz4e_cursor_skip(lz4e_cursor &c, uint32_t n)
{
	while (n) {
		const uint32_t room = c.v->len - c.done;

		if (n < room) {
			c.done += n;
			return;
		}
		n -= room;
		c.v++;
		c.done = 0;
	}
}

/* copy `len` bytes from linear memory `src` to the SG stream at cursor c */
template <int W>
LZ4E_DEV void lz4e_scatter(lz4e_cursor &c, const uint8_t *src, uint32_t len,
			   uint32_t l)
{
	while (len) {
		const uint32_t vlen = c.v->len;
		uint32_t take = vlen - c.done;

		if (take > len)
			take = len;
		lz4e_copy<W>(lz4e_cursor_addr(c), src, take, l);
		src += take;
		len -= take;
		c.done += take;
		if (c.done == vlen) {
			c.v++;
			c.done = 0;
		}
	}
}

/* copy `len` bytes from the SG stream at cursor c to linear memory `dst` */
template <int W>
LZ4E_DEV void lz4e_gather(uint8_t *dst, lz4e_cursor &c, uint32_t len,
			  uint32_t l)
{
	while (len) {
		const uint32_t vlen = c.v->len;
		uint32_t take = vlen - c.done;

		if (take > len)
			take = len;
		lz4e_copy<W>(dst, lz4e_cursor_addr(c), take, l);
		dst += take;
		len -= take;
		c.done += take;
		if (c.done == vlen) {
			c.v++;
			c.done = 0;
		}
	}
}

#endif /* LZ4E_COPY_CUH */
