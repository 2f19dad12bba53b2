/*
 * include/lz4e_bvec.h — userspace shim for the kernel scatter-gather types the
 * LZ4E API is written against: struct page, struct bio_vec, struct bvec_iter.
 *
 * Semantics follow Linux include/linux/bvec.h as the reference uses them
 * (doc/API.md:46-82 of the reference; lz4e/include/lz4e_defs.h:112-197 for the
 * iterator normal form):
 *   - a bio_vec names `bv_len` bytes starting `bv_offset` bytes into the page
 *     run that begins at `bv_page`; bv_offset + bv_len may exceed PAGE_SIZE
 *     (multi-page bvec), later pages are bv_page + k;
 *   - a bvec_iter selects a byte range of a bio_vec ARRAY: it starts
 *     `bi_bvec_done` bytes into bio_vec[bi_idx] and covers `bi_size` bytes;
 *   - iterators are normalised: bi_bvec_done < bio_vec[bi_idx].bv_len, a
 *     position at the very end of a bvec is (bi_idx + 1, 0).
 *
 * When this header is used inside a real kernel build the kernel's own
 * definitions are taken instead.
 */
#ifndef LZ4E_BVEC_H
#define LZ4E_BVEC_H

#ifdef __KERNEL__
#include <linux/bio.h>
#include <linux/bvec.h>
#else

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef PAGE_SHIFT
#define PAGE_SHIFT 12
#endif
#ifndef PAGE_SIZE
#define PAGE_SIZE (1UL << PAGE_SHIFT)
#endif
#ifndef BIO_MAX_VECS
#define BIO_MAX_VECS 256U
#endif

/* A userspace page IS its 4 KiB of memory, so `bv_page + k` steps whole pages
 * exactly like the kernel's struct page arithmetic does for a contiguous run. */
struct page {
	unsigned char mem[PAGE_SIZE];
};

static inline void *page_address(const struct page *page)
{
	return (void *)page->mem;
}

struct bio_vec {
	struct page *bv_page;
	unsigned int bv_len;
	unsigned int bv_offset;
};

typedef uint64_t sector_t;

struct bvec_iter {
	sector_t bi_sector;	   /* untouched by LZ4E */
	unsigned int bi_size;	   /* bytes covered from the current position */
	unsigned int bi_idx;	   /* current index into the bio_vec array */
	unsigned int bi_bvec_done; /* bytes already consumed in bio_vec[bi_idx] */
} __attribute__((packed, aligned(4)));

/* Minimal struct bio for the chunk layer (reference: doc/API.md:60-70,
 * lz4e_bdev/lz4e_chunk.c:139-149 reads bi_io_vec and bi_iter only). */
struct bio {
	struct bio_vec *bi_io_vec;
	unsigned short bi_vcnt;
	unsigned short bi_max_vecs;
	unsigned int bi_opf;
	struct bvec_iter bi_iter;
	void *bi_private;
};

/* bvec_iter_advance(): move `bytes` forward, keeping the normal form. */
static inline int lz4e_bvec_iter_advance(const struct bio_vec *bv,
					 struct bvec_iter *iter, unsigned bytes)
{
	unsigned int idx = iter->bi_idx;

	if (bytes > iter->bi_size) {
		iter->bi_size = 0;
		return 0;
	}
	iter->bi_size -= bytes;
	bytes += iter->bi_bvec_done;
	while (bytes && bytes >= bv[idx].bv_len) {
		bytes -= bv[idx].bv_len;
		idx++;
	}
	iter->bi_idx = idx;
	iter->bi_bvec_done = bytes;
	return 1;
}

#ifdef __cplusplus
}
#endif
#endif /* __KERNEL__ */
#endif /* LZ4E_BVEC_H */
