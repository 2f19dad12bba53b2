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
int lz4e_decompress_max_grid(int group, int stage, uint32_t stage_cap, int sm_count);
int lz4e_launch_decompress(const lz4e_decompress_args *a, int group, int stage,
			   int sm_count, cudaStream_t stream, int grid_limit,
			   int *grid_out);

#endif
