/*
 * include/lz4e_b200.h — the C ABI between the LZ4E host library and the CUDA
 * (sm_100a) kernels.  Plain pointers and sizes only.
 *
 * Three layers, lowest first:
 *
 *  1. DEVICE-LEVEL batch entry points (lz4e_b200_*_dev): every pointer is a
 *     device pointer, work is enqueued on a caller stream, nothing is copied.
 *     Blocks are described by device-side mirrors of the reference's own
 *     types: lz4e_b200_dev_bvec mirrors struct bio_vec (doc/API.md:46-52) with
 *     the page pointer already resolved to a device byte address, and
 *     lz4e_b200_dev_iter mirrors struct bvec_iter (doc/API.md:72-80) plus the
 *     index of the block's bio_vec[0] inside a shared table.  The kernels walk
 *     these exactly like LZ4E_fillBvIterSize / LZ4E_memcpy_from_sg do
 *     (lz4e/lz4e_compress.c:184-211, lz4e/include/lz4e_defs.h:366-402).
 *
 *  2. PAGE ARENAS: make a host `struct page *` resolvable on the device —
 *     pinned host pages with a device mirror (SURVEY.md section 8b, item iv).
 *
 *  3. HOST-LEVEL batch entry points (lz4e_b200_*_batch): arrays of the
 *     reference's call tuple {src, dst, srcIter, dstIter}; they translate the
 *     bio_vecs, move pages host<->device, run layer 1 and hand back the
 *     reference's return values.  LZ4E_compress_default (include/lz4e.h) is a
 *     batch of one.  This is what lz4e_chunk_compress_ext
 *     (lz4e_bdev/lz4e_chunk.c:139-159) would bind to.
 *
 * There is NO CPU fallback anywhere below this header.  Every function
 * returns LZ4E_B200_OK (0) or a negative LZ4E_B200_E* code and records a
 * message retrievable with lz4e_b200_strerror().
 */
#ifndef LZ4E_B200_H
#define LZ4E_B200_H

#include <stddef.h>
#include <stdint.h>

#include "lz4e_bvec.h"

#ifdef __cplusplus
extern "C" {
#endif

#define LZ4E_B200_OK 0
#define LZ4E_B200_EIO (-5)	/* CUDA runtime/driver error */
#define LZ4E_B200_ENOMEM (-12)
#define LZ4E_B200_EFAULT (-14)	/* page pointer outside every registered arena */
#define LZ4E_B200_ENODEV (-19)	/* no usable CUDA device: there is no CPU path */
#define LZ4E_B200_EINVAL (-22)
#define LZ4E_B200_ERANGE (-34)	/* a block broke a limit given to a *_dev call */

typedef struct lz4e_b200_ctx lz4e_b200_ctx;
typedef struct lz4e_b200_arena lz4e_b200_arena;

/* ---- device-side descriptor tables -------------------------------------- */

/* struct bio_vec with bv_page/bv_offset folded into a device byte address. */
struct lz4e_b200_dev_bvec {
	uint64_t addr; /* device address of page_address(bv_page) + bv_offset */
	uint32_t len;  /* raw bv_len (decides the hash-table address mode) */
	uint32_t rsvd;
};

/* struct bvec_iter for one block + where its bio_vec array starts. */
struct lz4e_b200_dev_iter {
	uint32_t first;	       /* index of the block's bio_vec[0] in the table */
	uint32_t bi_size;      /* input bytes / output capacity */
	uint32_t bi_idx;       /* as in struct bvec_iter */
	uint32_t bi_bvec_done; /* as in struct bvec_iter (mid-bvec start) */
};

/* Linear positions at which the reference leaves *srcIter / *dstIter
 * (lz4e/lz4e_compress.c:304-305, :528-529). */
struct lz4e_b200_dev_end {
	uint32_t src_pos;
	uint32_t dst_pos;
};

/* ---- context ------------------------------------------------------------- */
int lz4e_b200_device_count(void);
int lz4e_b200_ctx_create(int device, lz4e_b200_ctx **out);
/* waits for the device, releases the per-context host workspace and chunk
 * pool, and destroys every arena still registered with the context (their
 * handles are invalid afterwards) */
void lz4e_b200_ctx_destroy(lz4e_b200_ctx *ctx);
/* process-wide default context on $LZ4E_B200_DEVICE (default 0); NULL and a
 * message on stderr when no device is usable */
lz4e_b200_ctx *lz4e_b200_default_ctx(void);
const char *lz4e_b200_strerror(lz4e_b200_ctx *ctx);
/* kernel-selection knobs, the same ones the LZ4E_B200_* environment variables
 * set at context creation (tests and tuning; results never depend on them):
 * "dgroup" lanes per block of the decoder (1, 4, 8, 16; 0 = automatic),
 * "dstage" (-1/0/1), "dvariant" (-1 = automatic), "dorder", "wave_min",
 * "wave_serial", "lane_per_sm" (decoders), "gin_min", "gin_per_sm",
 * "gin_small", "tma", "p12" (compressor).  Returns LZ4E_B200_EINVAL for an
 * unknown key. */
int lz4e_b200_ctx_tune(lz4e_b200_ctx *ctx, const char *key, int value);
int lz4e_b200_ctx_device(lz4e_b200_ctx *ctx);
int lz4e_b200_ctx_sm_count(lz4e_b200_ctx *ctx);
/* kernels launched by this context so far (bench.py's gpu_launches) */
uint64_t lz4e_b200_kernel_launches(lz4e_b200_ctx *ctx);
/* wait for `stream` (NULL = the context's own stream) and report device-side
 * faults raised by earlier *_dev calls (LZ4E_B200_ERANGE ...) */
int lz4e_b200_sync(lz4e_b200_ctx *ctx, void *stream);

/* ---- layer 1: device-level batches --------------------------------------- */

/* promises a caller of a *_dev entry point may make (checked on the device,
 * a broken promise raises LZ4E_B200_ERANGE at the next lz4e_b200_sync) */
#define LZ4E_B200_DEV_NO_BYU16 1u   /* compress: no block selects the byU16 table
				     * (every block has a bv_len > 4096 or more
				     * than 16 bvecs): halves the table footprint */
#define LZ4E_B200_DEV_DST_LINEAR 2u /* every destination is one contiguous
				     * segment holding the whole capacity:
				     * decompress needs no scratch, compress
				     * writes in place without staging */
#define LZ4E_B200_DEV_SRC_LINEAR 4u /* decompress: every compressed block is
				     * one contiguous segment (no gather slot
				     * is reserved for the lane-per-block
				     * decoder) */

#define LZ4E_B200_DEV_DST_SEGS32 8u /* decompress: no destination has more than
				     * 32 segments under its iterator, so every
				     * block is decoded in place through the
				     * per-block segment map and no linear
				     * scratch slots are reserved */

/* largest block of one *_dev compress call: 16 MiB.  Up to 64 KiB the hash
 * table holds 16-bit positions; above, 32-bit ones (same number of entries as
 * lz4e_compress.c:48-57) and the block is read from a gather slot in global
 * memory (bios of 128 KiB ... 4 MiB, multi-page bvecs) */
#define LZ4E_B200_MAX_BLOCK (16u << 20)

/*
 * LZ4E_compress_default for `nblocks` independent blocks.
 *  src_bvecs/src_iters   source tables (device memory)
 *  dst_bvecs/dst_iters   destination tables; dst_iters[b].bi_size = capacity
 *  max_in_bytes          upper bound of src_iters[].bi_size (sizes the shared
 *                        memory staging; a larger block raises ERANGE)
 *  rets[b]               bytes written, or 0 exactly where the reference
 *                        returns 0
 *  ends[b]               optional (may be NULL) iterator end state
 * Output bytes are identical to the reference's for the same source bytes and
 * source segmentation; nothing beyond rets[b] bytes of a destination is
 * written.
 */
int lz4e_b200_compress_dev(lz4e_b200_ctx *ctx, void *stream,
			   const struct lz4e_b200_dev_bvec *src_bvecs,
			   const struct lz4e_b200_dev_iter *src_iters,
			   const struct lz4e_b200_dev_bvec *dst_bvecs,
			   const struct lz4e_b200_dev_iter *dst_iters,
			   uint32_t nblocks, uint32_t max_in_bytes,
			   uint32_t flags, int32_t *rets,
			   struct lz4e_b200_dev_end *ends);

/*
 * lz4e_b200_compress_dev for callers that must not allocate device memory
 * inside the call (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.
 *
 * Same arguments, same results (rets, ends, destination bytes) as
 * lz4e_b200_compress_dev.  It always runs the global-input compress kernel
 * (each warp gathers its block into a linear slot in global memory), and the
 * kernel's scratch — the work-queue counter and the input slots — lives in ONE
 * fixed array of LZ4E_B200_STATIC_SCRATCH_BYTES in the library's device image
 * (every GPU that loads the library holds it until the process exits) instead
 * of buffers the context grows with cudaMalloc.  The call itself makes no
 * cudaMalloc / cudaFree; the first call on a device looks the array's device
 * address up once (cudaGetSymbolAddress).
 *
 * Limits: max_in_bytes <= LZ4E_B200_STATIC_MAX_BLOCK, else LZ4E_B200_EINVAL
 * with a message in lz4e_b200_strerror(); any nblocks (the number of resident
 * warps is cut to the slots that fit the array: 127 slots at 64 KiB blocks, so
 * large blocks run with less parallelism than through lz4e_b200_compress_dev).
 * All calls — every context, every stream of a device — share the array: the
 * counter is cleared on the call's stream before the kernel, slots need no
 * clearing (each block overwrites what it reads), and two calls that overlap
 * on the GPU (different streams without an event between them) race on it and
 * produce garbage.  Serialise them.
 */
#define LZ4E_B200_STATIC_SCRATCH_BYTES (8u << 20)
#define LZ4E_B200_STATIC_MAX_BLOCK 65536u
int lz4e_b200_compress_dev_static(lz4e_b200_ctx *ctx, void *stream,
				  const struct lz4e_b200_dev_bvec *src_bvecs,
				  const struct lz4e_b200_dev_iter *src_iters,
				  const struct lz4e_b200_dev_bvec *dst_bvecs,
				  const struct lz4e_b200_dev_iter *dst_iters,
				  uint32_t nblocks, uint32_t max_in_bytes,
				  uint32_t flags, int32_t *rets,
				  struct lz4e_b200_dev_end *ends);

/*
 * LZ4E decompression (lz4e/lz4e_decompress.c:462-470 semantics on the
 * linearised bytes) with scatter-gather source and destination.
 *  src_iters[b].bi_size  compressed size,  dst_iters[b].bi_size  capacity
 *  rets[b]               decoded bytes, or the reference's negative error
 */
int lz4e_b200_decompress_dev(lz4e_b200_ctx *ctx, void *stream,
			     const struct lz4e_b200_dev_bvec *src_bvecs,
			     const struct lz4e_b200_dev_iter *src_iters,
			     const struct lz4e_b200_dev_bvec *dst_bvecs,
			     const struct lz4e_b200_dev_iter *dst_iters,
			     uint32_t nblocks, uint32_t max_in_bytes,
			     uint32_t max_out_bytes, uint32_t flags,
			     int32_t *rets);

/*
 * Compacted-output compression (SURVEY.md section 8b, item iii): all blocks
 * are written back to back into ONE destination SG stream: dst_bvecs[0 ..
 * dst_nbvecs), starting dst_bvec_done bytes into dst_bvecs[0], at most
 * dst_capacity bytes.  Blocks are compressed into context scratch (capacity
 * LZ4E_COMPRESSBOUND each, i.e. the reference's noLimit mode), a device prefix
 * scan over rets[] yields offsets[b] (exclusive; nblocks + 1 entries, the last
 * one is the total) and a pack kernel moves every block to its offset.  A
 * block that fails (ret 0) takes no space; running out of dst_capacity raises
 * LZ4E_B200_ERANGE at the next lz4e_b200_sync().
 */
int lz4e_b200_compress_compact_dev(lz4e_b200_ctx *ctx, void *stream,
				   const struct lz4e_b200_dev_bvec *src_bvecs,
				   const struct lz4e_b200_dev_iter *src_iters,
				   const struct lz4e_b200_dev_bvec *dst_bvecs,
				   uint32_t dst_nbvecs, uint32_t dst_bvec_done,
				   uint64_t dst_capacity, uint32_t nblocks,
				   uint32_t max_in_bytes, uint32_t flags,
				   int32_t *rets, uint64_t *offsets);

/* ---- raw memory helpers for the C host library --------------------------- */
void *lz4e_b200_dev_alloc(lz4e_b200_ctx *ctx, size_t bytes);
void lz4e_b200_dev_free(lz4e_b200_ctx *ctx, void *p);
void *lz4e_b200_host_alloc(lz4e_b200_ctx *ctx, size_t bytes); /* pinned */
void lz4e_b200_host_free(lz4e_b200_ctx *ctx, void *p);
/* asynchronous on `stream` (NULL = the context's stream) */
int lz4e_b200_copy_h2d(lz4e_b200_ctx *ctx, void *stream, void *dst_dev,
		       const void *src_host, size_t bytes);
int lz4e_b200_copy_d2h(lz4e_b200_ctx *ctx, void *stream, void *dst_host,
		       const void *src_dev, size_t bytes);

/*
 * Exact-byte transfers between arena host pages and their device mirror, done
 * by a kernel over PCIe instead of the copy engines: for block b the first
 * lens[b] bytes (lens == NULL: iters[b].bi_size bytes; lens[b] <= 0: none) of
 * the segment list (bvecs, iters[b]) — DEVICE addresses inside arenas of this
 * context — are copied device -> host (to_host != 0) or host -> device.
 * Segments that lie in no arena are skipped.  Nothing outside those bytes is
 * written on either side.  The host batches use it to bring back only the
 * bytes a compressor produced and to send up only the bytes a decompressor
 * will read.
 */
int lz4e_b200_xfer_dev(lz4e_b200_ctx *ctx, void *stream,
		       const struct lz4e_b200_dev_bvec *bvecs,
		       const struct lz4e_b200_dev_iter *iters, const int32_t *lens,
		       uint32_t nblocks, int to_host);
/* the same, leaving out the first `skip` bytes of every block (they travelled
 * another way, e.g. as the rows of a 2-D copy-engine transfer) */
int lz4e_b200_xfer_skip_dev(lz4e_b200_ctx *ctx, void *stream,
			    const struct lz4e_b200_dev_bvec *bvecs,
			    const struct lz4e_b200_dev_iter *iters, const int32_t *lens,
			    uint32_t nblocks, int to_host, uint32_t skip);
/* `height` rows of `width` bytes, rows `spitch` / `dpitch` bytes apart, pinned
 * host memory -> device, by copy engine */
int lz4e_b200_copy2d_h2d(lz4e_b200_ctx *ctx, void *stream, void *dst_dev, size_t dpitch,
			 const void *src_host, size_t spitch, size_t width, size_t height);
/* bytes moved host -> device and device -> host through this context so far
 * (copy-engine copies enqueued + transfer-kernel bytes completed); waits for
 * nothing but reads a device counter, so call it after lz4e_b200_sync() */
int lz4e_b200_traffic(lz4e_b200_ctx *ctx, uint64_t *h2d_bytes, uint64_t *d2h_bytes);

/* streams and events for callers that pipeline *_dev calls with copies (the
 * host-level batches do): opaque cudaStream_t / cudaEvent_t handles */
void *lz4e_b200_stream_create(lz4e_b200_ctx *ctx);
void lz4e_b200_stream_destroy(lz4e_b200_ctx *ctx, void *stream);
void *lz4e_b200_event_create(lz4e_b200_ctx *ctx);
void lz4e_b200_event_destroy(lz4e_b200_ctx *ctx, void *event);
/* an event that keeps a timestamp, and the time between two completed ones */
void *lz4e_b200_event_create_timed(lz4e_b200_ctx *ctx);
float lz4e_b200_event_elapsed_ms(lz4e_b200_ctx *ctx, void *from, void *to);
int lz4e_b200_event_record(lz4e_b200_ctx *ctx, void *event, void *stream);
/* work enqueued on `stream` after this call waits for `event` */
int lz4e_b200_stream_wait_event(lz4e_b200_ctx *ctx, void *stream, void *event);

/* ---- layer 2: page arenas ------------------------------------------------ */
int lz4e_b200_arena_create(lz4e_b200_ctx *ctx, size_t npages,
			   lz4e_b200_arena **out);
void lz4e_b200_arena_destroy(lz4e_b200_arena *arena);
struct page *lz4e_b200_arena_pages(lz4e_b200_arena *arena); /* pinned host */
void *lz4e_b200_arena_dev_base(lz4e_b200_arena *arena);	    /* device mirror */
size_t lz4e_b200_arena_npages(lz4e_b200_arena *arena);
int lz4e_b200_arena_upload(lz4e_b200_arena *arena, size_t first_page,
			   size_t npages);
int lz4e_b200_arena_download(lz4e_b200_arena *arena, size_t first_page,
			     size_t npages);
/* the context's arenas (at most `max` of them); returns how many were stored */
size_t lz4e_b200_arena_snapshot(lz4e_b200_ctx *ctx, lz4e_b200_arena **out,
				size_t max);
/* device address mirroring host address `p` if it lies in an arena of ctx,
 * else 0; *arena_out (optional) receives the arena */
uint64_t lz4e_b200_arena_translate(lz4e_b200_ctx *ctx, const void *p,
				   lz4e_b200_arena **arena_out);

/* ---- layer 3: host-level batches ----------------------------------------- */

/* One LZ4E call: the reference's argument tuple plus the return slot. */
struct lz4e_b200_req {
	const struct bio_vec *src;
	struct bio_vec *dst;
	struct bvec_iter *src_iter;
	struct bvec_iter *dst_iter;
	int ret;
	int rsvd;
};

#define LZ4E_B200_F_SRC_RESIDENT 1u /* source pages already current on device */
#define LZ4E_B200_F_DST_ON_DEVICE 2u /* leave results in the device mirror */
#define LZ4E_B200_F_KEEP_ITERS 4u   /* do not move the iterators */
/* No host byte other than the ones the codec produced may change (what a lone
 * LZ4E_compress_default call guarantees).  Compressed output always comes back
 * that way (rets[b] bytes per block through the transfer kernel); with this
 * flag a decompressor's output does too, instead of one copy-engine transfer
 * of the (densely tiled) destination range. */
#define LZ4E_B200_F_DST_PRESERVE 8u

/*
 * How pages travel (host/lz4e_host.c; measured split in
 * profiles/r2k_e2e_timeline.md):
 *  - pages in an arena, side tiles the byte range it spans (uncompressed
 *    blocks end to end): one copy-engine transfer of that range per chunk;
 *  - compressed output: exactly rets[b] bytes per block, written to the host
 *    pages by lz4e_xfer_kernel — never the LZ4E_COMPRESSBOUND slot;
 *  - sources that do NOT tile their range (compressed blocks sitting in their
 *    slots, scattered segments): the spanned range by copy engine when it is
 *    at least a quarter full, else exactly the bytes by the transfer kernel.
 *    The range copy refreshes the device mirror of the bytes IN BETWEEN from
 *    host memory; a caller that keeps device-only results
 *    (LZ4E_B200_F_DST_ON_DEVICE) in the same arena must not interleave them
 *    with another batch's sources, or sets LZ4E_B200_SPARSE_SRC=kernel;
 *  - pages anywhere else in host memory: packed into pinned staging by the
 *    context's host threads (sources) / copied back per segment (results).
 * A batch is cut into chunks that are pipelined over three streams; batches on
 * different contexts run concurrently, batches on one context take turns.
 * LZ4E_B200_TRACE=1 prints the host-side split per batch, =2 a device
 * timeline per chunk.
 */
int lz4e_b200_compress_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs,
			     size_t n, unsigned flags);
int lz4e_b200_decompress_batch(lz4e_b200_ctx *ctx, struct lz4e_b200_req *reqs,
			       size_t n, unsigned flags);

/* Contiguous range [first, first+count) of block indices owned by `rank`
 * out of `world` (shard by block index, no collective). */
void lz4e_b200_shard_range(uint64_t nblocks, uint32_t rank, uint32_t world,
			   uint64_t *first, uint64_t *count);

#ifdef __cplusplus
}
#endif
#endif /* LZ4E_B200_H */
