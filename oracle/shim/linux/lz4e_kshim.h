/*
 * oracle/shim/linux/lz4e_kshim.h — TEST INFRASTRUCTURE, not product code.
 *
 * Userspace stand-in for the handful of Linux kernel interfaces that the
 * reference's lz4e/lz4e_compress.c, lz4e/lz4e_decompress.c and
 * lz4e/include/lz4e_defs.h include.  It lets oracle/Makefile compile those
 * reference sources UNMODIFIED, from where they lie under /root/reference,
 * into oracle/_ref/liblz4e_ref.so.
 *
 * Everything here restates documented kernel semantics (Linux 6.17
 * include/linux/bvec.h, highmem.h, unaligned.h, bitops.h, minmax.h); none of
 * it comes from the reference repository.
 */
#ifndef LZ4E_ORACLE_KSHIM_H
#define LZ4E_ORACLE_KSHIM_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- compiler attributes ------------------------------------------------ */
#ifdef __always_inline
#undef __always_inline /* glibc's cdefs.h defines its own */
#endif
#define __always_inline inline __attribute__((__always_inline__))
#ifndef __packed
#define __packed __attribute__((__packed__))
#endif
#ifndef __aligned
#define __aligned(x) __attribute__((__aligned__(x)))
#endif
#define likely(x) __builtin_expect(!!(x), 1)
#define unlikely(x) __builtin_expect(!!(x), 0)

/* ---- architecture ------------------------------------------------------- */
#define CONFIG_64BIT 1
#define BITS_PER_LONG 64
#ifndef __LITTLE_ENDIAN
#define __LITTLE_ENDIAN 1234
#endif

/* ---- assertions --------------------------------------------------------- */
#define BUG_ON(cond)                                                        \
	do {                                                                \
		if (unlikely(cond)) {                                       \
			fprintf(stderr, "BUG_ON(%s) at %s:%d\n", #cond,     \
				__FILE__, __LINE__);                        \
			abort();                                            \
		}                                                           \
	} while (0)
#define WARN_ONCE(cond, ...) (unlikely(cond))
#define BUILD_BUG_ON(cond) ((void)sizeof(char[1 - 2 * !!(cond)]))

/* ---- module boilerplate ------------------------------------------------- */
#define EXPORT_SYMBOL(sym)
#define MODULE_AUTHOR(x)
#define MODULE_DESCRIPTION(x)
#define MODULE_LICENSE(x)

/* ---- min/max ------------------------------------------------------------ */
#define min(a, b)                                                           \
	({                                                                  \
		__typeof__(a) _a = (a);                                     \
		__typeof__(b) _b = (b);                                     \
		_a < _b ? _a : _b;                                          \
	})
#define min_t(type, a, b)                                                   \
	({                                                                  \
		type _a = (type)(a);                                        \
		type _b = (type)(b);                                        \
		_a < _b ? _a : _b;                                          \
	})

/* ---- bit operations ----------------------------------------------------- */
static inline unsigned long __ffs(unsigned long word)
{
	return (unsigned long)__builtin_ctzl(word);
}
static inline unsigned long __fls(unsigned long word)
{
	return (unsigned long)(BITS_PER_LONG - 1 - __builtin_clzl(word));
}

/* ---- unaligned access --------------------------------------------------- */
#define get_unaligned(ptr)                                                  \
	({                                                                  \
		const struct {                                              \
			__typeof__(*(ptr)) x;                               \
		} __packed *_gp = (const void *)(ptr);                      \
		_gp->x;                                                     \
	})
#define put_unaligned(val, ptr)                                             \
	do {                                                                \
		struct {                                                    \
			__typeof__(*(ptr)) x;                               \
		} __packed *_pp = (void *)(ptr);                            \
		_pp->x = (val);                                             \
	} while (0)
static inline uint16_t get_unaligned_le16(const void *p)
{
	const uint8_t *b = (const uint8_t *)p;
	return (uint16_t)(b[0] | (b[1] << 8));
}
static inline void put_unaligned_le16(uint16_t v, void *p)
{
	uint8_t *b = (uint8_t *)p;
	b[0] = (uint8_t)v;
	b[1] = (uint8_t)(v >> 8);
}

/* ---- pages -------------------------------------------------------------- */
#define PAGE_SHIFT 12
#define PAGE_SIZE (1UL << PAGE_SHIFT)
/* A page IS its 4 KiB of memory, so `bv_page + k` steps whole pages. */
struct page {
	unsigned char mem[PAGE_SIZE];
};
static inline void *page_address(const struct page *page)
{
	return (void *)page->mem;
}
static inline void memcpy_from_page(char *to, struct page *page, size_t offset,
				    size_t len)
{
	memcpy(to, (const char *)page_address(page) + offset, len);
}
static inline void memcpy_to_page(struct page *page, size_t offset,
				  const char *from, size_t len)
{
	memcpy((char *)page_address(page) + offset, from, len);
}

/* ---- bio_vec / bvec_iter (include/linux/bvec.h semantics) --------------- */
typedef uint64_t sector_t;
#define BIO_MAX_VECS 256U

struct bio_vec {
	struct page *bv_page;
	unsigned int bv_len;
	unsigned int bv_offset;
};

struct bvec_iter {
	sector_t bi_sector;
	unsigned int bi_size;
	unsigned int bi_idx;
	unsigned int bi_bvec_done;
} __packed __aligned(4);

#define __bvec_iter_bvec(bvec, iter) (&(bvec)[(iter).bi_idx])

/* multi-page view: the current bvec from the iterator position to its end */
#define mp_bvec_iter_page(bvec, iter) (__bvec_iter_bvec((bvec), (iter))->bv_page)
#define mp_bvec_iter_len(bvec, iter)                                        \
	min((iter).bi_size,                                                 \
	    __bvec_iter_bvec((bvec), (iter))->bv_len - (iter).bi_bvec_done)
#define mp_bvec_iter_offset(bvec, iter)                                     \
	(__bvec_iter_bvec((bvec), (iter))->bv_offset + (iter).bi_bvec_done)
#define mp_bvec_iter_page_idx(bvec, iter)                                   \
	(mp_bvec_iter_offset((bvec), (iter)) / PAGE_SIZE)

/* single-page view: clipped at the end of the page the position lies in */
#define bvec_iter_offset(bvec, iter)                                        \
	(mp_bvec_iter_offset((bvec), (iter)) % PAGE_SIZE)
#define bvec_iter_len(bvec, iter)                                           \
	min_t(unsigned, mp_bvec_iter_len((bvec), (iter)),                   \
	      PAGE_SIZE - bvec_iter_offset((bvec), (iter)))
#define bvec_iter_page(bvec, iter)                                          \
	(mp_bvec_iter_page((bvec), (iter)) +                                \
	 mp_bvec_iter_page_idx((bvec), (iter)))
#define bvec_iter_bvec(bvec, iter)                                          \
	((struct bio_vec){                                                  \
		.bv_page = bvec_iter_page((bvec), (iter)),                  \
		.bv_len = bvec_iter_len((bvec), (iter)),                    \
		.bv_offset = bvec_iter_offset((bvec), (iter)),              \
	})

static inline bool bvec_iter_advance(const struct bio_vec *bv,
				     struct bvec_iter *iter, unsigned bytes)
{
	unsigned int idx = iter->bi_idx;

	if (WARN_ONCE(bytes > iter->bi_size,
		      "Attempted to advance past end of bvec iter\n")) {
		iter->bi_size = 0;
		return false;
	}

	iter->bi_size -= bytes;
	bytes += iter->bi_bvec_done;

	while (bytes && bytes >= bv[idx].bv_len) {
		bytes -= bv[idx].bv_len;
		idx++;
	}

	iter->bi_idx = idx;
	iter->bi_bvec_done = bytes;
	return true;
}

/* advance inside the current bvec only (caller guarantees no overrun) */
static inline void bvec_iter_advance_single(const struct bio_vec *bv,
					    struct bvec_iter *iter,
					    unsigned int bytes)
{
	unsigned int done = iter->bi_bvec_done + bytes;

	if (done == bv[iter->bi_idx].bv_len) {
		done = 0;
		iter->bi_idx++;
	}
	iter->bi_bvec_done = done;
	iter->bi_size -= bytes;
}

#endif /* LZ4E_ORACLE_KSHIM_H */
