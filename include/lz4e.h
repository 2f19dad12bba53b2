/*
 * include/lz4e.h — public LZ4E API of the B200 implementation.
 *
 * Same names, argument meaning, sizing macros and error behaviour as the
 * reference's lz4e/include/lz4e.h:9-55 (documented in its doc/API.md:84-101),
 * so that a caller of the reference library — lz4e_chunk_compress_ext,
 * lz4e_bdev/lz4e_chunk.c:139-159 — can be relinked against liblz4e_b200.so
 * unchanged.  The work is done on a B200 by the CUDA kernels behind the C ABI in
 * include/lz4e_b200.h; there is no CPU implementation in this library: every
 * entry point fails (compress: 0, decompress: negative) with a message on
 * stderr if no CUDA device can be used.
 */
#ifndef LZ4E
#define LZ4E

#include "lz4e_bvec.h"

#ifdef __cplusplus
extern "C" {
#endif

#define LZ4E_NAME "lz4e"

#define LZ4E_ACCELERATION_DEFAULT 1 /* ref lz4e.h:9 */

/* ref lz4e.h:11-22: 16 KiB hash table + 1 KiB per-bvec sizes + 32 B misc */
#define LZ4E_MEMORY_USAGE 14
#define LZ4E_HASHLOG (LZ4E_MEMORY_USAGE - 2)
#define LZ4E_HASH_SIZE_U32 (1 << LZ4E_HASHLOG)
#define LZ4E_HASH_SIZE_U64 (LZ4E_HASH_SIZE_U32 >> 1)
#define LZ4E_BV_ITER_SIZE_U64 (BIO_MAX_VECS >> 1)
#define LZ4E_STREAMSIZE_U64 (LZ4E_HASH_SIZE_U64 + LZ4E_BV_ITER_SIZE_U64 + 4)
#define LZ4E_STREAMSIZE (LZ4E_STREAMSIZE_U64 * sizeof(unsigned long long))
#define LZ4E_MEM_COMPRESS LZ4E_STREAMSIZE /* 17 440 bytes */

/* ref lz4e.h:24-28 */
#define LZ4E_MAX_INPUT_SIZE 0x7E000000 /* 2 113 929 216 bytes */
#define LZ4E_COMPRESSBOUND(isize)                                           \
	((unsigned int)(isize) > (unsigned int)LZ4E_MAX_INPUT_SIZE          \
		 ? 0                                                        \
		 : (isize) + ((isize) / 255) + 16)

/*
 * ref lz4e.h:30-45: the stream state a caller may allocate instead of a raw
 * LZ4E_MEM_COMPRESS buffer.  Kept so that code declaring an LZ4E_stream_t (or
 * sizing wrkmem with sizeof) compiles and gets the same 17 440 bytes; this
 * library never reads the fields (per-bvec sizes and the hash table live in
 * device tables / shared memory), it only clears the buffer like
 * lz4e_compress.c:548 does.
 */
typedef struct {
	uint32_t bvIterSize[BIO_MAX_VECS];
	uint32_t hashTable[LZ4E_HASH_SIZE_U32];
	uint32_t currentOffset;
	uint32_t initCheck;
	const uint8_t *dictionary;
	uint8_t *bufferStart;
	uint32_t dictSize;
} LZ4E_stream_t_internal;
typedef union {
	unsigned long long table[LZ4E_STREAMSIZE_U64];
	LZ4E_stream_t_internal internal_donotuse;
} LZ4E_stream_t;

/* ref lz4e.h:53-55 */
#ifndef LZ4E_DISTANCE_MAX
#define LZ4E_DISTANCE_MAX 65535
#endif

/*
 * ref lz4e.h:47-48, lz4e/lz4e_compress.c:563-569.
 *
 * src / dst   bases of the source / destination bio_vec arrays
 * srcIter     start of the data inside src; bi_size = input length
 * dstIter     where to write inside dst; bi_size = output capacity
 * wrkmem      >= LZ4E_MEM_COMPRESS bytes owned by the caller; cleared like
 *             the reference does, otherwise unused (the hash table lives in
 *             the SM's shared memory)
 * returns     bytes written into dst, or 0 on failure (input larger than
 *             LZ4E_MAX_INPUT_SIZE, more than BIO_MAX_VECS source bvecs, or
 *             output capacity too small).  On success the iterators are left
 *             where the reference leaves them: *srcIter at the last match-search
 *             position, *dstIter at the first byte of the final literal run.
 *             On a 0 return caused by the output capacity the reference
 *             returns from inside its loop (lz4e_compress.c:358-363, :425-430,
 *             :509) with both iterators mid-way and a partial prefix in dst;
 *             the same state is produced here: *srcIter / *dstIter stand where
 *             the failing check found them and the destination holds the
 *             reference's bytes up to *dstIter.  Two things are unspecified,
 *             as they are garbage in the reference too: the token byte of the
 *             sequence in progress (never stored on these paths; it holds
 *             what the previous 4-byte length store or the caller left there)
 *             and the bytes behind *dstIter.  A block rejected up front
 *             (size, bvec count) leaves the iterators untouched, as there.
 */
int LZ4E_compress_default(const struct bio_vec *src, struct bio_vec *dst,
			  struct bvec_iter *srcIter, struct bvec_iter *dstIter,
			  void *wrkmem);

/*
 * ref lz4e.h:50-51, lz4e/lz4e_decompress.c:462-470: contiguous buffers.
 * returns decoded bytes (>= 0) or -(input offset where decoding stopped) - 1.
 */
int LZ4E_decompress_safe(const char *source, char *dest, int compressedSize,
			 int maxDecompressedSize);

/*
 * EXTENSION (not in the reference): the decompressor
 * with scatter-gather source and destination.  srcIter->bi_size is the
 * compressed size, dstIter->bi_size the output capacity.  Same return values
 * as LZ4E_decompress_safe on the linearised bytes.  Iterators are not moved.
 */
int LZ4E_decompress_safe_sg(const struct bio_vec *src, struct bio_vec *dst,
			    const struct bvec_iter *srcIter,
			    const struct bvec_iter *dstIter);

#ifdef __cplusplus
}
#endif
#endif /* LZ4E */
