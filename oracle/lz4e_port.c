/*
 * oracle/lz4e_port.c — TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement ("port") of the reference's LZ4E hot path in terms of LINEAR
 * byte positions instead of bvec iterators.  It exists so that
 *   (1) the equivalence "compressed bytes = f(linearised source, tableType)"
 *       that the CUDA kernels rely on is checked on the CPU against the
 *       reference's own object code (oracle/_ref/liblz4e_ref.so), and
 *   (2) a checker is available on machines where /root/reference is absent.
 *
 * Parity status: PINNED.  tests/test_oracle.py compares this file with
 * oracle/_ref (the reference sources compiled unmodified) on the reference's
 * three fixtures, on the SURVEY.md section 8(c) sanity values and on several
 * thousand fuzzed segmentations / malformed streams, and against the committed
 * vectors in tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
 * use this file.  The product path never links or loads it.
 *
 * Reference anchors (all under /root/reference):
 *   lz4e/lz4e_compress.c:184-211  LZ4E_fillBvIterSize   -> port_table_type()
 *   lz4e/lz4e_compress.c:48-96    hash log / hash4 / hash5 -> port_hash()
 *   lz4e/lz4e_compress.c:218-534  LZ4E_compress_generic -> port_compress_linear()
 *   lz4e/lz4e_compress.c:536-569  extState / default    -> LZ4E_compress_default()
 *   lz4e/lz4e_decompress.c:62-470 decompress_generic/safe -> LZ4E_decompress_safe()
 *   lz4e/include/lz4e_defs.h:587-636 LZ4E_count         -> port_lcp()
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <linux/bio.h>

enum { PORT_U16 = 1, PORT_U32 = 3, PORT_U64 = 7 }; /* lz4e_defs.h:689 */

#define PORT_MAX_INPUT 0x7E000000u /* lz4e.h:24 */
#define PORT_MAX_DISTANCE 65535u   /* lz4e.h:53-55 */
#define PORT_HASHLOG 12		   /* lz4e.h:11-12 */
#define PORT_TABLE_BYTES 16384	   /* hashTable[4096] u32, lz4e.h:13,35 */

static uint32_t rd32(const uint8_t *p)
{
	uint32_t v;
	memcpy(&v, p, 4);
	return v;
}

static uint64_t rd64(const uint8_t *p)
{
	uint64_t v;
	memcpy(&v, p, 8);
	return v;
}

/* lz4e_compress.c:48-96.  byU32 hashes 5 bytes to 12 bits, byU16 hashes 4
 * bytes to 13 bits, byU64 hashes 4 bytes to 11 bits. */
static uint32_t port_hash(const uint8_t *p, int tt)
{
	if (tt == PORT_U32)
		return (uint32_t)(((rd64(p) << 24) * 889523592379ULL) >>
				  (64 - PORT_HASHLOG));
	if (tt == PORT_U16)
		return (rd32(p) * 2654435761U) >> (32 - (PORT_HASHLOG + 1));
	return (rd32(p) * 2654435761U) >> (32 - (PORT_HASHLOG - 1));
}

/* exact longest common prefix, capped (lz4e_defs.h:587-636 computes the same
 * value with 8/4/2/1-byte steps) */
static uint32_t port_lcp(const uint8_t *a, const uint8_t *b, uint32_t limit)
{
	uint32_t n = 0;

	while (n < limit && a[n] == b[n])
		n++;
	return n;
}

struct port_out {
	uint8_t *buf;
	uint32_t pos;
};

static void put8(struct port_out *o, uint8_t v)
{
	o->buf[o->pos++] = v;
}

/*
 * Linear restatement of LZ4E_compress_generic (lz4e_compress.c:218-534).
 * d[0..n) source, out[0..cap) destination, tt the hash-table address mode.
 * Returns bytes written or 0.  *src_end / *dst_end receive the linear
 * positions the reference leaves *srcIter / *dstIter at on success.
 * `table` is 8192 zeroed u32 entries holding LINEAR positions
 * (0 == empty == position 0, equivalent to the (idx,off) encoding of
 * lz4e_defs.h:641-686 because that map is a bijection on positions).
 */
static int port_compress_linear(const uint8_t *d, uint32_t n, uint8_t *out,
				uint32_t cap, int tt, uint32_t *table,
				uint32_t *src_end, uint32_t *dst_end)
{
	const int limited = cap < (n + n / 255 + 16); /* lz4e_compress.c:553 */
	const uint32_t mflimit = n - 12;   /* :234, wraps for n<12 like the ref */
	const uint32_t matchlimit = n - 5; /* :235 */
	struct port_out o = { out, 0 };
	uint32_t ip = 0, anchor = 0, fwd_h, m = 0;
	uint32_t tokpos;
	uint8_t token;

	if (n > PORT_MAX_INPUT) /* :245-248 */
		return 0;
	if (n < 13) /* :268-271 */
		goto last_literals;

	table[port_hash(d, tt)] = 0; /* :280 */
	ip = 1;
	fwd_h = port_hash(d + ip, tt);

	for (;;) {
		/* find a match, :292-336 */
		{
			uint32_t fwd = ip, step = 1, nb = 1u << 6;

			for (;;) {
				uint32_t h = fwd_h;

				if (fwd + step > mflimit)
					goto last_literals;
				ip = fwd;
				fwd += step;
				step = nb++ >> 6;
				m = table[h];
				fwd_h = port_hash(d + fwd, tt);
				table[h] = ip;
				if (tt != PORT_U16 && m + PORT_MAX_DISTANCE < ip)
					continue;
				if (rd32(d + m) == rd32(d + ip))
					break;
			}
		}
		/* catch up, :339-349 */
		while (ip > anchor && m > 0 && d[ip - 1] == d[m - 1]) {
			ip--;
			m--;
		}
		/* literals, :352-382 */
		{
			uint32_t lit = ip - anchor, i;

			tokpos = o.pos++;
			if (limited && o.pos + lit + 8 + lit / 255 > cap)
				return 0;
			if (lit >= 15) {
				uint32_t len = lit - 15;

				token = 15 << 4;
				for (; len >= 255; len -= 255)
					put8(&o, 255);
				put8(&o, (uint8_t)len);
			} else {
				token = (uint8_t)(lit << 4);
			}
			for (i = 0; i < lit; i++)
				o.buf[o.pos + i] = d[anchor + i];
			o.pos += lit;
		}
next_match:
		/* offset, :385-387 */
		put8(&o, (uint8_t)(ip - m));
		put8(&o, (uint8_t)((ip - m) >> 8));
		/* match length, :420-450 */
		{
			uint32_t mc;

			ip += 4;
			m += 4;
			mc = port_lcp(d + ip, d + m, matchlimit - ip);
			ip += mc;
			if (limited && o.pos + 6 + (mc >> 8) > cap)
				return 0;
			if (mc >= 15) {
				token += 15;
				mc -= 15;
				for (; mc >= 255; mc -= 255)
					put8(&o, 255);
				put8(&o, (uint8_t)mc);
			} else {
				token += (uint8_t)mc;
			}
			o.buf[tokpos] = token;
		}
		anchor = ip;
		if (ip > mflimit) /* :456 */
			break;
		table[port_hash(d + ip - 2, tt)] = ip - 2; /* :461-464 */
		/* test next position, :467-493 */
		{
			uint32_t h = port_hash(d + ip, tt);

			m = table[h];
			table[h] = ip;
			if (m + PORT_MAX_DISTANCE >= ip &&
			    rd32(d + m) == rd32(d + ip)) {
				token = 0;
				tokpos = o.pos++;
				goto next_match;
			}
		}
		ip++;
		fwd_h = port_hash(d + ip, tt);
	}

last_literals:
	/* :500-530 */
	{
		uint32_t run = n - anchor;

		if (limited && o.pos + run + 1 + (run + 255 - 15) / 255 > cap)
			return 0;
		if (run >= 15) {
			uint32_t acc = run - 15;

			put8(&o, 15 << 4);
			for (; acc >= 255; acc -= 255)
				put8(&o, 255);
			put8(&o, (uint8_t)acc);
		} else {
			put8(&o, (uint8_t)(run << 4));
		}
		*dst_end = o.pos; /* dstIter is not moved past the last run */
		memcpy(o.buf + o.pos, d + anchor, run);
		o.pos += run;
	}
	*src_end = ip;
	return (int)o.pos;
}

/* lz4e_compress.c:184-211.  Returns 0 when more than BIO_MAX_VECS bvecs lie
 * under the iterator.  Looks at the RAW bv_len of every covered bvec. */
static int port_table_type(const struct bio_vec *bv, struct bvec_iter it,
			   int *tt)
{
	unsigned int start_idx = it.bi_idx;

	*tt = PORT_U16;
	while (it.bi_size) {
		unsigned int i = it.bi_idx - start_idx;
		unsigned int len = bv[it.bi_idx].bv_len;
		unsigned int take = len - it.bi_bvec_done;

		if (i >= BIO_MAX_VECS)
			return 0;
		if (i >= 16 || len > 4096)
			*tt |= PORT_U32;
		if (len > (1u << 24))
			*tt |= PORT_U64;
		if (take > it.bi_size)
			take = it.bi_size;
		bvec_iter_advance_single(bv, &it, take);
	}
	return 1;
}

static void port_gather(uint8_t *to, const struct bio_vec *bv,
			struct bvec_iter it, uint32_t n)
{
	while (n) {
		const struct bio_vec *v = &bv[it.bi_idx];
		unsigned int take = v->bv_len - it.bi_bvec_done;

		if (take > n)
			take = n;
		memcpy(to,
		       (const uint8_t *)page_address(v->bv_page) + v->bv_offset +
			       it.bi_bvec_done,
		       take);
		bvec_iter_advance_single(bv, &it, take);
		to += take;
		n -= take;
	}
}

static void port_scatter(struct bio_vec *bv, struct bvec_iter it,
			 const uint8_t *from, uint32_t n)
{
	while (n) {
		struct bio_vec *v = &bv[it.bi_idx];
		unsigned int take = v->bv_len - it.bi_bvec_done;

		if (take > n)
			take = n;
		memcpy((uint8_t *)page_address(v->bv_page) + v->bv_offset +
			       it.bi_bvec_done,
		       from, take);
		bvec_iter_advance_single(bv, &it, take);
		from += take;
		n -= take;
	}
}

/* lz4e_compress.c:536-569 */
int LZ4E_compress_default(const struct bio_vec *src, struct bio_vec *dst,
			  struct bvec_iter *srcIter, struct bvec_iter *dstIter,
			  void *wrkmem)
{
	const uint32_t n = srcIter->bi_size;
	const uint32_t cap = dstIter->bi_size;
	uint32_t bound, src_end = 0, dst_end = 0;
	uint8_t *lin, *out;
	uint32_t *table;
	int tt = PORT_U16, ret;

	memset(wrkmem, 0, 17440); /* :548 */
	if (n > PORT_MAX_INPUT)
		return 0;
	if (n >= 13 && !port_table_type(src, *srcIter, &tt)) /* :268-277 */
		return 0;
	bound = n + n / 255 + 16;
	lin = (uint8_t *)malloc((size_t)n + 16);
	out = (uint8_t *)malloc((size_t)bound + 16);
	if (!lin || !out)
		abort();
	port_gather(lin, src, *srcIter, n);
	/* 8192 linear positions (the largest table, byU16) as plain u32; the
	 * reference packs the same entries into wrkmem's 16 KiB, lz4e.h:33-35 */
	table = (uint32_t *)calloc(8192, sizeof(uint32_t));
	if (!table)
		abort();
	ret = port_compress_linear(lin, n, out, cap < bound ? cap : bound, tt,
				   table, &src_end, &dst_end);
	free(table);
	if (cap >= bound && ret == 0 && n <= PORT_MAX_INPUT)
		abort(); /* noLimit mode cannot fail */
	if (ret > 0) {
		port_scatter(dst, *dstIter, out, (uint32_t)ret);
		bvec_iter_advance(src, srcIter, src_end);
		bvec_iter_advance(dst, dstIter, dst_end);
	}
	free(lin);
	free(out);
	return ret;
}

/*
 * Restatement of LZ4E_decompress_generic specialised the way
 * LZ4E_decompress_safe instantiates it (endOnInputSize, decode_full_block,
 * noDict, lowPrefix == dest), lz4e_decompress.c:62-470, on integer offsets.
 * Error value = -(input offset reached) - 1 (:458-459).
 */
int LZ4E_decompress_safe(const char *source, char *dest, int compressedSize,
			 int maxDecompressedSize)
{
	const uint8_t *src = (const uint8_t *)source;
	uint8_t *dst = (uint8_t *)dest;
	const int64_t iend = compressedSize, oend = maxDecompressedSize;
	int64_t ip = 0, op = 0;

	if (oend == 0) /* :113-114 */
		return (iend == 1 && src[0] == 0) ? 0 : -1;
	if (iend == 0) /* :119-120 */
		return -1;

	for (;;) {
		const unsigned token = src[ip++];
		uint64_t length = token >> 4;
		int64_t offset, match, i;

		/* two-stage shortcut, :150-191 */
		if (length != 15 && ip < iend - 16 && op <= oend - 32) {
			for (i = 0; i < (int64_t)length; i++)
				dst[op + i] = src[ip + i];
			op += (int64_t)length;
			ip += (int64_t)length;
			length = token & 15;
			offset = src[ip] | (src[ip + 1] << 8);
			ip += 2;
			match = op - offset;
			if (length != 15 && offset >= 8 && match >= 0) {
				for (i = 0; i < (int64_t)length + 4; i++)
					dst[op + i] = dst[match + i];
				op += (int64_t)length + 4;
				continue;
			}
			goto copy_match;
		}

		/* literal length, :194-220 */
		if (length == 15) {
			unsigned s;

			if (ip >= iend - 15)
				goto error;
			do {
				s = src[ip++];
				length += s;
			} while (ip < iend - 15 && s == 255);
		}

		/* literals, :222-277 */
		if (op + (int64_t)length > oend - 12 ||
		    ip + (int64_t)length > iend - 8) {
			if (ip + (int64_t)length != iend ||
			    op + (int64_t)length > oend)
				goto error;
			memmove(dst + op, src + ip, length);
			op += (int64_t)length;
			break;
		}
		memcpy(dst + op, src + ip, length);
		ip += (int64_t)length;
		op += (int64_t)length;

		offset = src[ip] | (src[ip + 1] << 8); /* :280-283 */
		ip += 2;
		match = op - offset;
		length = token & 15;

copy_match:
		if (match < 0) /* :299-302 */
			goto error;
		if (length == 15) { /* :316-334 */
			unsigned s;

			do {
				s = src[ip++];
				if (ip > iend - 5)
					goto error;
				length += s;
			} while (s == 255);
		}
		length += 4;
		if (op + (int64_t)length > oend - 12 && /* :420-431 */
		    op + (int64_t)length > oend - 5)
			goto error;
		if (offset == 0) { /* LZ4_write32(op, offset) at :309-314 */
			memset(dst + op, 0, length);
		} else {
			for (i = 0; i < (int64_t)length; i++)
				dst[op + i] = dst[match + i];
		}
		op += (int64_t)length;
	}
	return (int)op;

error:
	return (int)(-ip) - 1;
}
