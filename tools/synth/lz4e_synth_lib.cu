/*
 * tools/synth/lz4e_synth_lib.cu — BENCHMARK/TEST DATA GENERATOR, not product
 * code: fills blocks with the synthetic contents of lz4e_synth.h (SURVEY.md
 * section 8d, config 2) on the host or straight into device-resident
 * scatter-gather blocks.  Built as tools/synth/liblz4e_synth.so; neither the
 * product library (lz4-sgori_b200/liblz4e_b200.so) nor the oracle link it.
 *
 * The device entry point takes the same descriptor tables the product's
 * device-level ABI takes (include/lz4e_b200.h: lz4e_b200_dev_bvec /
 * lz4e_b200_dev_iter) but runs on its own: plain CUDA runtime calls on the
 * current device, its own scratch, its own two kernels.
 */
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <thread>
#include <vector>

#include "../../include/lz4e_b200.h"
#include "lz4e_synth.h"

/* one thread writes one linear block */
__global__ void synth_linear_kernel(uint8_t *lin, uint32_t stride,
				    const struct lz4e_b200_dev_iter *iters,
				    uint32_t nblocks, uint32_t first_index)
{
	const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;

	if (b >= nblocks)
		return;
	const uint32_t n = iters[b].bi_size < stride ? iters[b].bi_size : stride;

	lz4e_synth_block(lin + (size_t)b * stride, n, (first_index + b) & 3u, first_index + b);
}

/* linear slots -> SG blocks, one warp per block, byte per lane (set-up code) */
__global__ void synth_scatter_kernel(const uint8_t *lin, uint32_t stride,
				     const struct lz4e_b200_dev_bvec *bvecs,
				     const struct lz4e_b200_dev_iter *iters, uint32_t nblocks)
{
	const uint32_t lane = threadIdx.x & 31;

	for (uint32_t b = blockIdx.x * 4 + (threadIdx.x >> 5); b < nblocks; b += gridDim.x * 4) {
		const struct lz4e_b200_dev_iter it = iters[b];
		const struct lz4e_b200_dev_bvec *v = bvecs + it.first + it.bi_idx;
		const uint8_t *s = lin + (size_t)b * stride;
		uint32_t done = it.bi_bvec_done;
		uint32_t left = it.bi_size < stride ? it.bi_size : stride;

		while (left) {
			uint32_t take = v->len - done;
			uint8_t *g = (uint8_t *)(uintptr_t)(v->addr + done);

			if (take > left)
				take = left;
			for (uint32_t i = lane; i < take; i += 32)
				g[i] = s[i];
			s += take;
			left -= take;
			done = 0;
			v++;
		}
	}
}

extern "C" {

void lz4e_synth_fill_host(uint8_t *buf, uint32_t n, uint32_t cls, uint32_t seed)
{
	lz4e_synth_block(buf, n, cls, seed);
}

/* nblocks blocks of n bytes at base + b * stride: class (first + b) & 3, seed
 * first + b; nthreads host threads */
void lz4e_synth_fill_host_blocks(uint8_t *base, uint64_t stride, uint32_t n, uint64_t nblocks,
				 uint64_t first, int nthreads)
{
	if (nthreads < 1)
		nthreads = 1;
	std::vector<std::thread> th;

	for (int t = 0; t < nthreads; t++)
		th.emplace_back([=]() {
			for (uint64_t b = (uint64_t)t; b < nblocks; b += (uint64_t)nthreads)
				lz4e_synth_block(base + b * stride, n, (uint32_t)((first + b) & 3u),
						 (uint32_t)(first + b));
		});
	for (auto &x : th)
		x.join();
}

/* Fill device-resident SG blocks on the current device of `device`; blocks
 * larger than max_bytes are truncated to it.  Synchronous.  Returns 0 or -1. */
int lz4e_synth_fill_dev(int device, const struct lz4e_b200_dev_bvec *bvecs,
			const struct lz4e_b200_dev_iter *iters, uint32_t nblocks,
			uint32_t first_index, uint32_t max_bytes)
{
	const uint32_t stride = (max_bytes + 255u) & ~255u;
	uint32_t chunk = (uint32_t)(((size_t)512 << 20) / (stride ? stride : 256u));
	uint8_t *lin = nullptr;
	int rc = 0;

	if (chunk < 1)
		chunk = 1;
	if (chunk > nblocks)
		chunk = nblocks;
	if (nblocks == 0 || stride == 0)
		return 0;
	if (cudaSetDevice(device) != cudaSuccess ||
	    cudaMalloc((void **)&lin, (size_t)chunk * stride) != cudaSuccess) {
		fprintf(stderr, "lz4e_synth: device set-up failed: %s\n",
			cudaGetErrorString(cudaGetLastError()));
		return -1;
	}
	for (uint32_t b0 = 0; b0 < nblocks && !rc; b0 += chunk) {
		const uint32_t nb = nblocks - b0 < chunk ? nblocks - b0 : chunk;
		unsigned grid = (nb + 3) / 4;

		if (grid > 148u * 16u)
			grid = 148u * 16u;
		synth_linear_kernel<<<(nb + 63) / 64, 64>>>(lin, stride, iters + b0, nb, first_index + b0);
		synth_scatter_kernel<<<grid, 128>>>(lin, stride, bvecs, iters + b0, nb);
		if (cudaDeviceSynchronize() != cudaSuccess) {
			fprintf(stderr, "lz4e_synth: fill failed: %s\n",
				cudaGetErrorString(cudaGetLastError()));
			rc = -1;
		}
	}
	cudaFree(lin);
	return rc;
}

} /* extern "C" */
