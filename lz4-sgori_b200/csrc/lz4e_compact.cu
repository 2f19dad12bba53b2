/*
 * lz4e_compact.cu — K3: compacted-output compression (extension iii of
 * SURVEY.md section 8b; the reference has no batching, its chunk layer gives
 * every bio its own LZ4E_COMPRESSBOUND-sized buffer,
 * lz4e_bdev/lz4e_chunk.c:56-97).
 *
 *   1. lz4e_slots_kernel   describes one bound-sized linear slot per block
 *   2. lz4e_compress_kernel (K2) compresses every block into its slot
 *   3. lz4e_scan_kernel    exclusive prefix sums: compressed sizes -> offsets,
 *                          destination bvec lengths -> stream positions
 *   4. lz4e_pack_kernel    one warp per block copies slot -> destination SG
 *                          stream at its offset (binary search over the bvec
 *                          prefix, then the cursor copy of lz4e_copy.cuh)
 */
#include <cuda_runtime.h>

#include "../../include/lz4e_b200.h"
#include "lz4e_compress_body.cuh"
#include "lz4e_kernels.h"

extern "C" {
uint8_t *lz4e_ctx_scratch(lz4e_b200_ctx *ctx, void *stream, size_t bytes);
uint32_t *lz4e_ctx_faults(lz4e_b200_ctx *ctx);
void *lz4e_ctx_stream(lz4e_b200_ctx *ctx, void *stream);
void lz4e_ctx_count_launches(lz4e_b200_ctx *ctx, unsigned k);
}

__global__ void lz4e_slots_kernel(const struct lz4e_b200_dev_iter *src_iters,
				  uint8_t *slots, uint32_t stride,
				  struct lz4e_b200_dev_bvec *bv,
				  struct lz4e_b200_dev_iter *it, uint32_t nblocks)
{
	const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;

	if (b >= nblocks)
		return;
	const uint32_t n = src_iters[b].bi_size;

	bv[b].addr = (uint64_t)(uintptr_t)(slots + (size_t)b * stride);
	bv[b].len = stride;
	bv[b].rsvd = 0;
	it[b].first = b;
	it[b].bi_size = n + n / 255u + 16u; /* LZ4E_COMPRESSBOUND, lz4e.h:25-28 */
	it[b].bi_idx = 0;
	it[b].bi_bvec_done = 0;
}

/* out[i] = sum of value(j) for j < i, i in [0, n]; single CTA of 1024 threads.
 * MODE 0: value = max(rets[j], 0); MODE 1: value = bvecs[j].len */
template <int MODE>
__global__ void __launch_bounds__(1024)
lz4e_scan_kernel(const void *in, uint32_t n, uint64_t *out)
{
	__shared__ uint64_t part[1024];
	const uint32_t t = threadIdx.x;
	const uint32_t chunk = (n + 1023u) / 1024u;
	const uint32_t lo = t * chunk;
	const uint32_t hi = lo + chunk < n ? lo + chunk : n;
	uint64_t sum = 0;

	for (uint32_t i = lo; i < hi; i++) {
		if (MODE == 0) {
			const int32_t r = ((const int32_t *)in)[i];
			sum += r > 0 ? (uint32_t)r : 0u;
		} else {
			sum += ((const struct lz4e_b200_dev_bvec *)in)[i].len;
		}
	}
	part[t] = sum;
	__syncthreads();
	for (uint32_t d = 1; d < 1024; d <<= 1) { /* Hillis-Steele inclusive scan */
		uint64_t v = t >= d ? part[t - d] : 0;

		__syncthreads();
		part[t] += v;
		__syncthreads();
	}
	uint64_t run = t ? part[t - 1] : 0;

	for (uint32_t i = lo; i < hi; i++) {
		out[i] = run;
		if (MODE == 0) {
			const int32_t r = ((const int32_t *)in)[i];
			run += r > 0 ? (uint32_t)r : 0u;
		} else {
			run += ((const struct lz4e_b200_dev_bvec *)in)[i].len;
		}
	}
	if (t == 1023)
		out[n] = part[1023];
}

__global__ void __launch_bounds__(128)
lz4e_pack_kernel(const uint8_t *slots, uint32_t stride, const int32_t *rets,
		 const uint64_t *offsets, const struct lz4e_b200_dev_bvec *dst_bvecs,
		 const uint64_t *dst_prefix, uint32_t dst_nbvecs, uint64_t dst_skip,
		 uint64_t dst_cap, uint32_t nblocks, uint32_t *faults)
{
	const uint32_t lane = threadIdx.x & 31;

	for (uint32_t b = blockIdx.x * 4 + (threadIdx.x >> 5); b < nblocks;
	     b += gridDim.x * 4) {
		const int32_t r = rets[b];

		if (r <= 0)
			continue;
		const uint64_t off = offsets[b];

		if (off + (uint32_t)r > dst_cap) {
			if (lane == 0)
				atomicOr(faults, LZ4E_FAULT_RANGE);
			continue;
		}
		/* stream position -> bvec: last j with dst_prefix[j] <= pos */
		const uint64_t pos = off + dst_skip;
		uint32_t lo = 0, hi = dst_nbvecs;

		while (hi - lo > 1) {
			const uint32_t mid = (lo + hi) >> 1;

			if (dst_prefix[mid] <= pos)
				lo = mid;
			else
				hi = mid;
		}
		lz4e_cursor c;

		c.v = dst_bvecs + lo;
		c.done = (uint32_t)(pos - dst_prefix[lo]);
		lz4e_scatter<32>(c, slots + (size_t)b * stride, (uint32_t)r, lane);
	}
}

extern "C" int lz4e_b200_compress_compact_dev(
	lz4e_b200_ctx *ctx, void *stream, const struct lz4e_b200_dev_bvec *src_bvecs,
	const struct lz4e_b200_dev_iter *src_iters,
	const struct lz4e_b200_dev_bvec *dst_bvecs, uint32_t dst_nbvecs,
	uint32_t dst_bvec_done, uint64_t dst_capacity, uint32_t nblocks,
	uint32_t max_in_bytes, uint32_t flags, int32_t *rets, uint64_t *offsets)
{
	if (!ctx)
		return LZ4E_B200_ENODEV;
	if (nblocks == 0)
		return LZ4E_B200_OK;
	cudaSetDevice(lz4e_b200_ctx_device(ctx));
	cudaStream_t s = (cudaStream_t)lz4e_ctx_stream(ctx, stream);
	const uint32_t bound = max_in_bytes + max_in_bytes / 255u + 16u;
	const uint32_t stride = (bound + 255u) & ~255u;
	const size_t slot_bytes = (size_t)nblocks * stride;
	const size_t bv_bytes = ((size_t)nblocks * sizeof(lz4e_b200_dev_bvec) + 255) & ~(size_t)255;
	const size_t it_bytes = ((size_t)nblocks * sizeof(lz4e_b200_dev_iter) + 255) & ~(size_t)255;
	const size_t pf_bytes = (((size_t)dst_nbvecs + 1) * 8 + 255) & ~(size_t)255;
	uint8_t *scr = lz4e_ctx_scratch(ctx, stream, slot_bytes + bv_bytes + it_bytes + pf_bytes);

	if (!scr)
		return LZ4E_B200_ENOMEM;
	lz4e_b200_dev_bvec *sbv = (lz4e_b200_dev_bvec *)(scr + slot_bytes);
	lz4e_b200_dev_iter *sit = (lz4e_b200_dev_iter *)(scr + slot_bytes + bv_bytes);
	uint64_t *prefix = (uint64_t *)(scr + slot_bytes + bv_bytes + it_bytes);

	lz4e_slots_kernel<<<(nblocks + 255) / 256, 256, 0, s>>>(src_iters, scr, stride, sbv,
								sit, nblocks);
	int rc = lz4e_b200_compress_dev(ctx, (void *)s, src_bvecs, src_iters, sbv, sit, nblocks,
					max_in_bytes, flags | LZ4E_B200_DEV_DST_LINEAR, rets,
					nullptr);
	if (rc)
		return rc;
	lz4e_scan_kernel<0><<<1, 1024, 0, s>>>(rets, nblocks, offsets);
	lz4e_scan_kernel<1><<<1, 1024, 0, s>>>(dst_bvecs, dst_nbvecs, prefix);
	long long grid = ((long long)nblocks + 3) / 4;
	const long long maxg = (long long)lz4e_b200_ctx_sm_count(ctx) * 16;

	if (grid > maxg)
		grid = maxg;
	lz4e_pack_kernel<<<(unsigned)grid, 128, 0, s>>>(
		scr, stride, rets, offsets, dst_bvecs, prefix, dst_nbvecs, dst_bvec_done,
		dst_capacity, nblocks, lz4e_ctx_faults(ctx));
	lz4e_ctx_count_launches(ctx, 4);
	return cudaGetLastError() == cudaSuccess ? LZ4E_B200_OK : LZ4E_B200_EIO;
}
