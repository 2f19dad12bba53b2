/*
 * lz4e_xfer.cu — exact-byte transfers between an arena's pinned host pages and
 * their device mirror, done by the SMs over PCIe (the host pages are
 * cudaHostAlloc memory, directly addressable from the device under UVA).
 *
 * The host-level batches (host/lz4e_host.c) used to move whole byte RANGES of
 * an arena with the copy engines: a compressed 4 KiB block that takes 1.7 KB
 * came back as its whole 4 224-byte destination slot, and went up again as a
 * whole slot for decompression.  Here one warp walks the block's own segment
 * list and moves exactly the bytes that count — rets[b] produced bytes
 * (device -> host, "push") or bi_size source bytes (host -> device, "pull") —
 * with 16-byte vectors, four in flight per lane.  Nothing outside those bytes
 * is touched on either side, which also removes every "bulk range crosses
 * somebody else's pages" hazard of the range copies.
 *
 * What it stands in for on the reference side: nothing on the codec path — the
 * reference works on host memory in place (lz4e/include/lz4e_defs.h:366-402
 * touches bio pages directly); this is the price of the device boundary.
 */
#include <cuda_runtime.h>

#include "lz4e_kernels.h"
#include "lz4e_copy.cuh"

__global__ void __launch_bounds__(256) lz4e_xfer_kernel(const lz4e_xfer_args a)
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
	unsigned long long moved = 0;

	for (uint32_t b = warp; b < a.nblocks; b += nwarps) {
		const struct lz4e_b200_dev_iter it = a.iters[b];
		uint32_t len = it.bi_size;

		if (a.lens) {
			const int32_t r = a.lens[b];

			len = r > 0 ? ((uint32_t)r < len ? (uint32_t)r : len) : 0u;
		}
		const struct lz4e_b200_dev_bvec *v = a.bvecs + it.first + it.bi_idx;
		uint32_t done = it.bi_bvec_done;
		int last = 0;

		if (a.skip) { /* the head of the block travelled another way */
			uint32_t sk = a.skip < len ? a.skip : len;

			len -= sk;
			while (sk) {
				const uint32_t room = v->len - done;

				if (sk < room) {
					done += sk;
					break;
				}
				sk -= room;
				done = 0;
				v++;
			}
		}

		while (len) {
			const uint32_t vlen = v->len;
			uint32_t take = vlen - done;
			const uint64_t dev = v->addr + done;
			int k = -1;

			if (take > len)
				take = len;
			if (dev >= a.ar[last].lo && dev + take <= a.ar[last].hi) {
				k = last;
			} else {
				for (int i = 0; i < a.narenas; i++)
					if (dev >= a.ar[i].lo && dev + take <= a.ar[i].hi)
						k = i;
			}
			if (k >= 0) { /* segments outside every arena are bounced by the host */
				uint8_t *d = (uint8_t *)(uintptr_t)dev;
				uint8_t *h = (uint8_t *)(uintptr_t)(dev + a.ar[k].delta);

				last = k;
				if (a.to_host)
					lz4e_copy<32>(h, d, take, lane);
				else
					lz4e_copy<32>(d, h, take, lane);
				moved += take;
			}
			len -= take;
			done = 0;
			v++;
		}
	}
	if (a.moved && lane == 0 && moved)
		atomicAdd(a.moved, moved);
}

int lz4e_launch_xfer(const lz4e_xfer_args *a, int sm_count, cudaStream_t stream)
{
	/* 8 warps per CTA; enough CTAs to keep a few hundred KB in flight over
	 * PCIe, never more warps than blocks */
	const uint32_t want = (a->nblocks + 7u) / 8u;
	uint32_t grid = (uint32_t)sm_count * 2u;

	if (grid > want)
		grid = want;
	if (grid == 0)
		return 0;
	lz4e_xfer_kernel<<<grid, 256, 0, stream>>>(*a);
	return cudaGetLastError() != cudaSuccess;
}
