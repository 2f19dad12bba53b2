/*
 * lz4e_order.cu — cost-ordered work queue for the decoders.
 *
 * Blocks of one batch differ a lot in decoding cost: the number of sequences
 * follows the compressed size (a sequence is at least 3 bytes), except for
 * incompressible blocks, which are one literal run.  Two consequences on a
 * GPU: (1) a long block that starts late ends late while the rest of the
 * machine idles, (2) the groups of one warp advance in lockstep, so a warp
 * whose groups hold blocks of different kinds executes the union of their code
 * paths at every step.  Both go away when the blocks are handed out in order of
 * decreasing estimated cost: a counting sort of the block indices by
 * compressed size relative to the output capacity (64 buckets, incompressible
 * blocks last).  Results do not depend on the order; it only changes which
 * block a group takes next.
 *
 * Two small kernels: per-CTA shared-memory histograms -> global histogram;
 * then every CTA reserves a range per bucket and writes its block indices.
 */
#include <cuda_runtime.h>

#include "../../include/lz4e_b200.h"
#include "lz4e_kernels.h"

#define ORDER_TILE 1024u
#define ORDER_THREADS 256u

static __device__ __forceinline__ uint32_t order_key(const struct lz4e_b200_dev_iter &s,
						     const struct lz4e_b200_dev_iter *dp,
						     uint32_t b, uint32_t by_size)
{
	const uint32_t c = s.bi_size;

	if (by_size) {
		/* compression: cost follows the input size; largest first */
		const uint32_t k = (uint32_t)(((uint64_t)c * (LZ4E_ORDER_BUCKETS - 1u)) / by_size);

		return LZ4E_ORDER_BUCKETS - 1u - (k < LZ4E_ORDER_BUCKETS ? k : LZ4E_ORDER_BUCKETS - 1u);
	}
	const uint32_t cap = dp[b].bi_size;

	if (cap == 0 || c >= cap)
		return LZ4E_ORDER_BUCKETS - 1u; /* incompressible or malformed: cheap */
	/* larger compressed size first */
	return LZ4E_ORDER_BUCKETS - 2u - (uint32_t)(((uint64_t)c * (LZ4E_ORDER_BUCKETS - 1u)) / cap);
}

__global__ void __launch_bounds__(ORDER_THREADS)
lz4e_order_count_kernel(const struct lz4e_b200_dev_iter *src_iters,
			const struct lz4e_b200_dev_iter *dst_iters, uint32_t n, uint32_t *hist,
			uint32_t by_size)
{
	__shared__ uint32_t h[LZ4E_ORDER_BUCKETS];

	if (threadIdx.x < LZ4E_ORDER_BUCKETS)
		h[threadIdx.x] = 0;
	__syncthreads();
	const uint32_t base = blockIdx.x * ORDER_TILE;

	for (uint32_t i = threadIdx.x; i < ORDER_TILE; i += ORDER_THREADS) {
		const uint32_t b = base + i;

		if (b < n)
			atomicAdd(&h[order_key(src_iters[b], dst_iters, b, by_size)], 1u);
	}
	__syncthreads();
	if (threadIdx.x < LZ4E_ORDER_BUCKETS && h[threadIdx.x])
		atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void __launch_bounds__(ORDER_THREADS)
lz4e_order_place_kernel(const struct lz4e_b200_dev_iter *src_iters,
			const struct lz4e_b200_dev_iter *dst_iters, uint32_t n,
			const uint32_t *hist, uint32_t *cursor, uint32_t *order, uint32_t by_size)
{
	__shared__ uint32_t h[LZ4E_ORDER_BUCKETS];    /* this tile's counts */
	__shared__ uint32_t at[LZ4E_ORDER_BUCKETS];   /* where this tile's bucket starts */
	uint32_t key[ORDER_TILE / ORDER_THREADS], rank[ORDER_TILE / ORDER_THREADS];
	const uint32_t base = blockIdx.x * ORDER_TILE;

	if (threadIdx.x < LZ4E_ORDER_BUCKETS)
		h[threadIdx.x] = 0;
	__syncthreads();
#pragma unroll
	for (uint32_t k = 0; k < ORDER_TILE / ORDER_THREADS; k++) {
		const uint32_t b = base + k * ORDER_THREADS + threadIdx.x;

		key[k] = 0;
		rank[k] = 0;
		if (b < n) {
			key[k] = order_key(src_iters[b], dst_iters, b, by_size);
			rank[k] = atomicAdd(&h[key[k]], 1u);
		}
	}
	__syncthreads();
	if (threadIdx.x < LZ4E_ORDER_BUCKETS) {
		uint32_t prefix = 0;

		for (uint32_t j = 0; j < threadIdx.x; j++)
			prefix += hist[j];
		at[threadIdx.x] = prefix + (h[threadIdx.x] ? atomicAdd(&cursor[threadIdx.x],
								       h[threadIdx.x])
							   : 0u);
	}
	__syncthreads();
#pragma unroll
	for (uint32_t k = 0; k < ORDER_TILE / ORDER_THREADS; k++) {
		const uint32_t b = base + k * ORDER_THREADS + threadIdx.x;

		if (b < n)
			order[at[key[k]] + rank[k]] = b;
	}
}

/* ctl: 2 * LZ4E_ORDER_BUCKETS words of device memory (histogram + cursors) */
int lz4e_launch_order(const struct lz4e_b200_dev_iter *src_iters,
		      const struct lz4e_b200_dev_iter *dst_iters, uint32_t n, uint32_t *ctl,
		      uint32_t *order, cudaStream_t stream, uint32_t by_size)
{
	const unsigned grid = (n + ORDER_TILE - 1u) / ORDER_TILE;

	if (cudaMemsetAsync(ctl, 0, 2 * LZ4E_ORDER_BUCKETS * sizeof(uint32_t), stream) != cudaSuccess)
		return -1;
	lz4e_order_count_kernel<<<grid, ORDER_THREADS, 0, stream>>>(src_iters, dst_iters, n, ctl,
								    by_size);
	lz4e_order_place_kernel<<<grid, ORDER_THREADS, 0, stream>>>(
		src_iters, dst_iters, n, ctl, ctl + LZ4E_ORDER_BUCKETS, order, by_size);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
