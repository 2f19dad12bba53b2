/*
 * include/lz4e_chunk.h — userspace form of the lz4e_bdev chunk layer and of its
 * request path, on top of the GPU LZ4E library.
 *
 * Reference interfaces mirrored here (names, argument meaning, return values):
 *   lz4e_bdev/include/lz4e_chunk.h:16-46   struct lz4e_buffer, struct lz4e_chunk,
 *                                          lz4e_chunk_alloc/free/compress/
 *                                          compress_ext/decompress,
 *                                          lz4e_buf_copy_from_bio
 *   lz4e_bdev/lz4e_chunk.c:19-159          their behaviour: 0 on success, -EIO when
 *                                          the codec reports failure, a size
 *                                          mismatch after decompression is fatal
 *   lz4e_bdev/lz4e_req.c:144-213           WRITE = chunk_alloc -> new bio over the
 *                                          destination buffer -> compress_ext ->
 *                                          decompress -> store the DECOMPRESSED
 *                                          bytes; :53-75 READ = pass-through
 *   lz4e_bdev/include/lz4e_stats.h:16-21,  reqs_total / reqs_failed / vec_count /
 *   lz4e_bdev/lz4e_stats.c:39-52           data_in_bytes per direction
 *
 * What is different, and why: the reference handles one bio per call on the CPU.
 * A GPU wants many independent blocks per launch, so every operation also exists
 * in a *_batch form that takes an array of chunks (or bios) and issues ONE
 * lz4e_b200_compress_batch / lz4e_b200_decompress_batch for all of them; the
 * single-chunk functions are batches of one.  Chunk buffers come from a pool of
 * pinned pages with a device mirror (lz4e_b200 page arena), which is what makes
 * `virt_to_page(data)` (lz4e_req.c:125) resolvable on the device.  There is no
 * CPU codec behind any of this: without a CUDA device the calls fail with -EIO.
 */
#ifndef LZ4E_CHUNK_H
#define LZ4E_CHUNK_H

#include <stddef.h>
#include <stdint.h>

#include "lz4e_b200.h"
#include "lz4e_bvec.h"

#ifdef __cplusplus
extern "C" {
#endif

/* contiguous data in memory (lz4e_chunk.h:16-21 of the reference) */
struct lz4e_buffer {
	struct bio *bio;
	char *data;
	int data_size;
	int buf_size;
};

/* data to be compressed (lz4e_chunk.h:24-28 of the reference) */
struct lz4e_chunk {
	struct lz4e_buffer src_buf;
	struct lz4e_buffer dst_buf;
	void *wrkmem;
};

/* context the chunk layer runs on (default: lz4e_b200_default_ctx()); call
 * before the first allocation.  pool_pages = 0 keeps the default pool size
 * ($LZ4E_CHUNK_POOL_MB, 1024 MiB). */
int lz4e_chunk_setup(lz4e_b200_ctx *ctx, size_t pool_pages);
/* Several GPUs from one process: every context has its own pool (created by
 * lz4e_chunk_setup(ctx, ...) or on first use), and a thread chooses the context
 * its chunk and request-path calls run on with lz4e_chunk_bind(ctx) (NULL:
 * back to the default).  One submitter thread per GPU is the intended shape;
 * a batch runs on the context whose pool holds its chunks. */
void lz4e_chunk_bind(lz4e_b200_ctx *ctx);
/* release the pool (all chunks must have been freed) */
void lz4e_chunk_teardown(void);

void lz4e_buf_copy_from_bio(struct lz4e_buffer *dst, struct bio *src);
struct lz4e_chunk *lz4e_chunk_alloc(int src_size);
void lz4e_chunk_free(struct lz4e_chunk *chunk);
/* src_buf.data -> dst_buf.data, contiguous.  NOTE: the reference calls the
 * kernel's stock LZ4_compress_default here (lz4e_bdev/lz4e_chunk.c:106-107);
 * this build runs the LZ4E codec on a one-segment (multi-page bvec, byU32)
 * description of the two buffers.  Both emit valid LZ4 blocks that either
 * decoder accepts, but the BYTES differ from the reference's for buffers above
 * 4 KiB (different hash-table mode); only round trips are guaranteed here.
 * Byte identity holds for lz4e_chunk_compress_ext, the LZ4E path. */
int lz4e_chunk_compress(struct lz4e_chunk *chunk);
/* dst_buf.data -> src_buf.data; the decoded size must equal src_buf.buf_size */
int lz4e_chunk_decompress(struct lz4e_chunk *chunk);
/* src_buf.bio -> dst_buf.bio with LZ4E_compress_default semantics */
int lz4e_chunk_compress_ext(struct lz4e_chunk *chunk);

/* batched forms: status[i] = 0 or -EIO per chunk; the return value is 0 when
 * the batch ran (individual chunks may still have failed) */
int lz4e_chunk_compress_batch(struct lz4e_chunk **chunks, size_t n, int *status);
int lz4e_chunk_decompress_batch(struct lz4e_chunk **chunks, size_t n, int *status);
int lz4e_chunk_compress_ext_batch(struct lz4e_chunk **chunks, size_t n, int *status);

/* wrap a chunk buffer into page bio_vecs, as lz4e_add_buf_to_bio
 * (lz4e_req.c:115-142) does with bio_add_page; returns 0 or -EAGAIN when the
 * bio has too few bi_max_vecs */
int lz4e_add_buf_to_bio(struct bio *bio, struct lz4e_buffer *buf);

/* ---- request path (row f2) ------------------------------------------------ */

struct lz4e_stats {
	uint64_t reqs_total;
	uint64_t reqs_failed;
	uint64_t vec_count;
	uint64_t data_in_bytes;
};

/* stand-in for the proxy device: a RAM "underlying device" of `capacity`
 * bytes plus read and write statistics */
struct lz4e_dev;

struct lz4e_dev *lz4e_dev_create(uint64_t capacity);
void lz4e_dev_destroy(struct lz4e_dev *dev);
/*
 * lz4e_write_req_init for `n` write bios at once: every bio is compressed as
 * ONE block from its own bio_vecs into a fresh chunk, decompressed again,
 * checked, and the decompressed bytes are stored at bi_iter.bi_sector * 512.
 * status[i] is 0 or a negative errno (-EIO codec failure, -ENOMEM, -ERANGE
 * outside the device).
 */
int lz4e_dev_submit_writes(struct lz4e_dev *dev, struct bio **bios, size_t n,
			   int *status);
/* READ: bytes of the underlying device into the bio's pages */
int lz4e_dev_submit_reads(struct lz4e_dev *dev, struct bio **bios, size_t n,
			  int *status);
void lz4e_dev_stats(struct lz4e_dev *dev, int write, struct lz4e_stats *out);
void lz4e_dev_stats_reset(struct lz4e_dev *dev);
/* bytes the codec produced for the writes so far (compressed size) */
uint64_t lz4e_dev_compressed_bytes(struct lz4e_dev *dev);
/* order-independent digest of WHAT the codec produced: the wrapping sum, over
 * all successful writes so far, of lz4e_fnv1a64() (a word-wise FNV-1a variant) of each bio's compressed
 * bytes — lets a harness compare the bytes of every bio with the reference's
 * without keeping them */
uint64_t lz4e_dev_compressed_digest(struct lz4e_dev *dev);
uint64_t lz4e_fnv1a64(const void *p, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* LZ4E_CHUNK_H */
