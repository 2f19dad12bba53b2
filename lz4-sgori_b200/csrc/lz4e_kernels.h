/* lz4e_kernels.h — launch helpers implemented in lz4e_kernels.cu */
#ifndef LZ4E_KERNELS_H
#define LZ4E_KERNELS_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lz4e_b200.h"

#define LZ4E_D_WARPS 4 /* warps per decompression CTA */

struct lz4e_compress_args;
struct lz4e_decompress_args;

int lz4e_compress_plan(lz4e_compress_args *a, int direct, int gin, int sm_count,
		       int max_per_sm, int *grid_out, size_t *smem_out);
int lz4e_launch_compress(const lz4e_compress_args *a, int direct, int gin, int grid,
			 size_t smem, cudaStream_t stream);
/* decoder variants: the round-1 group decoder (lz4e_decompress_body.cuh), the
 * group decoder of lz4e_decompress_group.cuh without / with pipelined matches */
#define LZ4E_DV_OLD 0
#define LZ4E_DV_GROUP 1
#define LZ4E_DV_GROUP_PIPE 2
#define LZ4E_DV_WAVE 3 /* lz4e_decompress_wave.cuh: one warp per block */

int lz4e_decompress_max_grid(int group, int stage, uint32_t stage_cap, int sm_count,
			     int variant);
int lz4e_launch_decompress(const lz4e_decompress_args *a, int group, int stage,
			   int variant, int sm_count, cudaStream_t stream, int grid_limit,
			   int *grid_out);

/* cost-ordered work queue (lz4e_order.cu): order[] receives the block indices
 * by decreasing estimated decoding cost; ctl = 2 * LZ4E_ORDER_BUCKETS words */
#define LZ4E_ORDER_BUCKETS 64u
int lz4e_launch_order(const struct lz4e_b200_dev_iter *src_iters,
		      const struct lz4e_b200_dev_iter *dst_iters, uint32_t n, uint32_t *ctl,
		      uint32_t *order, cudaStream_t stream, uint32_t by_size = 0);
/* by_size != 0: order by source size alone (largest first, by_size = the
 * largest size + 1): the compressor's cost estimate; dst_iters is not read */


/* exact-byte host <-> device transfers by the SMs (lz4e_xfer.cu) */
#define LZ4E_XFER_MAX_ARENAS 16
struct lz4e_xfer_args {
	const struct lz4e_b200_dev_bvec *bvecs; /* DEVICE addresses */
	const struct lz4e_b200_dev_iter *iters;
	const int32_t *lens; /* bytes per block (<= 0: none); NULL: bi_size */
	unsigned long long *moved; /* byte counter, may be NULL */
	uint32_t nblocks;
	uint32_t skip; /* leave out the first `skip` bytes of every block */
	int to_host;
	int narenas;
	struct {
		uint64_t lo, hi; /* device range of the arena's mirror */
		uint64_t delta;	 /* host address = device address + delta (mod 2^64) */
	} ar[LZ4E_XFER_MAX_ARENAS];
};
int lz4e_launch_xfer(const lz4e_xfer_args *a, int sm_count, cudaStream_t stream);

#endif
