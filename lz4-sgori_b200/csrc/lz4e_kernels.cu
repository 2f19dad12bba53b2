/*
 * lz4e_kernels.cu — __global__ entry points (sm_100a) around the warp bodies
 * and their launch helpers.  The C ABI in lz4e_abi.cu calls the
 * lz4e_launch_* functions declared in lz4e_kernels.h.
 */
#include <cuda_runtime.h>

#include "lz4e_compress_body.cuh"
#include "lz4e_decompress_body.cuh"
#include "lz4e_decompress_lane.cuh"
#include "lz4e_decompress_group.cuh"
#include "lz4e_decompress_wave.cuh"
#include "lz4e_kernels.h"

extern __shared__ __align__(16) uint8_t lz4e_smem[];

/* ---- K2: compression, one warp per block in flight; a.warps warps per CTA,
 * each with its own a.warp_smem bytes of shared memory ---------------------- */
#define LZ4E_C_MAX_WARPS 24
template <bool DIRECT, bool GIN, bool TMA = false>
__global__ void __launch_bounds__(LZ4E_C_MAX_WARPS * 32)
lz4e_compress_kernel(const lz4e_compress_args a)
{
	const uint32_t w = threadIdx.x >> 5;
	uint32_t lane = threadIdx.x & 31;

	/* opaque to the compiler, which otherwise re-reads SR_TID.X (S2R, ~25
	 * cycles) in about a hundred places instead of keeping the lane index */
	asm volatile("" : "+r"(lane));
	/* the same for this warp's shared-memory base: left to itself the compiler
	 * rebuilds the address of lz4e_smem (S2R SR_CgaCtaId, MOV, LEA) in front of
	 * about 75 shared-memory loads, on the encoder's dependent chains */
	uint32_t soff = (uint32_t)__cvta_generic_to_shared(lz4e_smem) + w * a.warp_smem;

	/* (the 32-bit shared-space address is what is hidden: a pointer that came
	 * out of an asm statement would lose its "is shared memory" property) */
	asm volatile("" : "+r"(soff));
	uint8_t *sm = (uint8_t *)__cvta_shared_to_generic((size_t)soff);

	lz4e_compress_warp<DIRECT, GIN, TMA>(a, sm, lane, blockIdx.x * a.warps + w);
}

/* ---- K4: decompression, LZ4E_D_WARPS warps per CTA, 32/GS blocks per warp */
template <int GS, bool STAGE>
__global__ void __launch_bounds__(LZ4E_D_WARPS * 32)
lz4e_decompress_kernel(const lz4e_decompress_args a)
{
	const uint32_t warp = threadIdx.x >> 5;
	const uint32_t groups = 32 / GS;
	uint8_t *sm = lz4e_smem + (size_t)warp * groups * (a.stage_cap + LZ4E_D_PAD);
	const uint32_t slot0 = (blockIdx.x * LZ4E_D_WARPS + warp) * groups;

	lz4e_decompress_warp_impl<GS, STAGE>(a, sm, slot0, threadIdx.x & 31);
}

/* ---- K4G: the group decoder of lz4e_decompress_group.cuh: in-place scattered
 * output through a per-group segment map (or staged in shared memory), parse
 * window in registers, pipelined short matches (PIPE) ------------------------ */
template <int GS, bool STAGE, bool PIPE>
__global__ void __launch_bounds__(LZ4E_D_WARPS * 32)
lz4e_decompress_group_kernel(const lz4e_decompress_args a)
{
	const uint32_t warp = threadIdx.x >> 5;
	constexpr uint32_t groups = 32 / GS;
	const size_t per_group = STAGE ? (size_t)(a.stage_cap + LZ4E_D_PAD) : sizeof(lz4e_omap);
	uint8_t *sm = lz4e_smem + (size_t)warp * groups * per_group;
	const uint32_t slot0 = (blockIdx.x * LZ4E_D_WARPS + warp) * groups;

	lz4e_decompress_groups<GS, STAGE, PIPE>(a, sm, slot0, threadIdx.x & 31);
}

/* ---- K4L: lane-per-block decompression, 32 blocks in flight per warp ------ */
__global__ void __launch_bounds__(LZ4E_D_WARPS * 32)
lz4e_decompress_lane_kernel(const lz4e_decompress_args a)
{
	__shared__ lz4e_lane_smem sm[LZ4E_D_WARPS];
	const uint32_t warp = threadIdx.x >> 5;
	const uint32_t slot0 = (blockIdx.x * LZ4E_D_WARPS + warp) * 32u;

	lz4e_decompress_lanes(a, &sm[warp], slot0, threadIdx.x & 31);
}

/* ---- K4W: the wave decoder of lz4e_decompress_wave.cuh: one warp per block,
 * 32 sequences per step.  Ten CTAs per SM (what their shared memory allows)
 * instead of the eight that 64 registers gave: 48 registers, one word spilled,
 * 64 KiB blocks over 16 segments 244 -> 253 GB/s ------------------------------ */
#ifndef LZ4E_W_MINB
#define LZ4E_W_MINB 10
#endif
__global__ void __launch_bounds__(LZ4E_D_WARPS * 32, LZ4E_W_MINB)
lz4e_decompress_wave_kernel(const lz4e_decompress_args a)
{
	__shared__ lz4e_wave_smem sm[LZ4E_D_WARPS];
	const uint32_t warp = threadIdx.x >> 5;

	uint32_t lane = threadIdx.x & 31;

	/* opaque to the compiler: under the 48-register cap it otherwise re-reads
	 * SR_TID.X (S2R, ~25 cycles) twice per step of the copy loop instead of
	 * keeping the lane index (253 -> 259 GB/s at 64 KiB x 16 segments) */
	asm volatile("" : "+r"(lane));
	lz4e_decompress_wave(a, &sm[warp], blockIdx.x * LZ4E_D_WARPS + warp, lane);
}

/* ------------------------------------------------------------------------- */

static int lz4e_set_smem(const void *fn, size_t bytes)
{
	if (bytes <= 48 * 1024)
		return 0;
	return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
				    (int)bytes) == cudaSuccess
		       ? 0
		       : -1;
}

typedef void (*lz4e_compress_fn)(const lz4e_compress_args);

/* tma: the shared-memory kernels with 1-D TMA bulk staging of the input (its
 * own instantiation: merely compiling the cp.async.bulk / mbarrier path into
 * the plain kernel costs it 10 %, profiles/r2o_tma_staging.md) */
static lz4e_compress_fn compress_fn(int direct, int gin, int tma = 0)
{
	if (!gin && tma)
		return direct ? lz4e_compress_kernel<true, false, true>
			      : lz4e_compress_kernel<false, false, true>;
	if (gin)
		return direct ? lz4e_compress_kernel<true, true> : lz4e_compress_kernel<false, true>;
	return direct ? lz4e_compress_kernel<true, false> : lz4e_compress_kernel<false, false>;
}

/*
 * Grid and shared-memory size of a compress launch: one persistent one-warp
 * CTA per block in flight, as many per SM as shared memory allows, at most
 * max_per_sm when that is > 0 (the GIN kernels read their input through L1,
 * which shares the SM's 256 KB with shared memory: fewer, better cached warps
 * can beat more warps).
 */
int lz4e_compress_plan(lz4e_compress_args *a, int direct, int gin, int sm_count,
		       int max_per_sm, int *grid_out, size_t *smem_out)
{
	size_t smem = lz4e_compress_smem_bytes(a->table_bytes, gin ? 0u : a->in_cap, a->out_cap);
	int per_sm = 0;
	lz4e_compress_fn fn = compress_fn(direct, gin, (int)a->tma);

	smem = (smem + 15) & ~(size_t)15;
	a->warps = 1;
	a->warp_smem = (uint32_t)smem;
	if (max_per_sm > 0) {
		/* ask for enough shared memory that only max_per_sm CTAs fit */
		const size_t want = ((size_t)227 * 1024 / (size_t)max_per_sm - 1024) & ~(size_t)127;

		if (want > smem)
			smem = want;
	}
	if (lz4e_set_smem((const void *)fn, smem))
		return -1;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 32, smem) !=
		    cudaSuccess ||
	    per_sm < 1)
		return -1;
	if (max_per_sm <= 0) {
		/* every CTA pays 1 KB of reserved shared memory: ONE CTA of W warps
		 * per SM can hold more warps than W one-warp CTAs (4 KiB blocks
		 * written in place: 11 instead of 10) */
		const size_t cta_max = (size_t)227 * 1024;
		int w = (int)(cta_max / a->warp_smem);

		if (w > LZ4E_C_MAX_WARPS)
			w = LZ4E_C_MAX_WARPS;
		if (w > per_sm && (long long)a->nblocks >= (long long)w * sm_count) {
			int one = 0;
			const size_t big = (size_t)w * a->warp_smem;

			if (lz4e_set_smem((const void *)fn, big) == 0 &&
			    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&one, fn, w * 32, big) ==
				    cudaSuccess &&
			    one >= 1) {
				a->warps = (uint32_t)w;
				*grid_out = sm_count * one;
				*smem_out = big;
				return 0;
			}
			(void)cudaGetLastError();
			if (lz4e_set_smem((const void *)fn, smem))
				return -1;
		}
	}
	long long grid = (long long)per_sm * sm_count;

	if (grid > (long long)a->nblocks)
		grid = a->nblocks;
	if (grid < 1)
		grid = 1;
	*grid_out = (int)grid;
	*smem_out = smem;
	return 0;
}

int lz4e_launch_compress(const lz4e_compress_args *a, int direct, int gin, int grid,
			 size_t smem, cudaStream_t stream)
{
	lz4e_compress_fn fn = compress_fn(direct, gin, (int)a->tma);

	fn<<<(unsigned)grid, 32u * a->warps, smem, stream>>>(*a);
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int GS, bool STAGE>
static int launch_decompress_t(const lz4e_decompress_args *a, int sm_count,
			       cudaStream_t stream, int grid_limit, int *grid_out)
{
	const size_t smem =
		STAGE ? (size_t)LZ4E_D_WARPS * (32 / GS) * (a->stage_cap + LZ4E_D_PAD) : 0;
	int per_sm = 0;
	auto fn = lz4e_decompress_kernel<GS, STAGE>;

	if (lz4e_set_smem((const void *)fn, smem))
		return -1;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
		    &per_sm, fn, LZ4E_D_WARPS * 32, smem) != cudaSuccess ||
	    per_sm < 1)
		return -1;
	long long grid = (long long)per_sm * sm_count;
	const long long want =
		((long long)a->nblocks + LZ4E_D_WARPS * (32 / GS) - 1) /
		(LZ4E_D_WARPS * (32 / GS));

	if (grid > want)
		grid = want;
	if (grid_limit > 0 && grid > grid_limit)
		grid = grid_limit;
	if (grid < 1)
		grid = 1;
	fn<<<(unsigned)grid, LZ4E_D_WARPS * 32, smem, stream>>>(*a);
	if (grid_out)
		*grid_out = (int)grid;
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

static int launch_decompress_lanes(const lz4e_decompress_args *a, int sm_count,
				   cudaStream_t stream, int grid_limit, int *grid_out)
{
	int per_sm = 0;

	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
		    &per_sm, lz4e_decompress_lane_kernel, LZ4E_D_WARPS * 32, 0) != cudaSuccess ||
	    per_sm < 1)
		return -1;
	long long grid = (long long)per_sm * sm_count;
	const long long want =
		((long long)a->nblocks + LZ4E_D_WARPS * 32 - 1) / (LZ4E_D_WARPS * 32);

	if (grid > want)
		grid = want;
	if (grid_limit > 0 && grid > grid_limit)
		grid = grid_limit;
	if (grid < 1)
		grid = 1;
	lz4e_decompress_lane_kernel<<<(unsigned)grid, LZ4E_D_WARPS * 32, 0, stream>>>(*a);
	if (grid_out)
		*grid_out = (int)grid;
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

static int wave_per_sm(void)
{
	int per_sm = 0;

	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
		    &per_sm, lz4e_decompress_wave_kernel, LZ4E_D_WARPS * 32, 0) != cudaSuccess ||
	    per_sm < 1)
		return -1;
	return per_sm;
}

static int launch_decompress_wave(const lz4e_decompress_args *a, int sm_count,
				  cudaStream_t stream, int grid_limit, int *grid_out)
{
	const int per_sm = wave_per_sm();

	if (per_sm < 1)
		return -1;
	long long grid = (long long)per_sm * sm_count;
	const long long want = ((long long)a->nblocks + LZ4E_D_WARPS - 1) / LZ4E_D_WARPS;

	if (grid > want)
		grid = want;
	if (grid_limit > 0 && grid > grid_limit)
		grid = grid_limit;
	if (grid < 1)
		grid = 1;
	lz4e_decompress_wave_kernel<<<(unsigned)grid, LZ4E_D_WARPS * 32, 0, stream>>>(*a);
	if (grid_out)
		*grid_out = (int)grid;
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

typedef void (*lz4e_dgroup_fn)(const lz4e_decompress_args);

template <bool STAGE, bool PIPE>
static lz4e_dgroup_fn dgroup_fn_gs(int group)
{
	switch (group) {
	case 4:
		return lz4e_decompress_group_kernel<4, STAGE, PIPE>;
	case 16:
		return lz4e_decompress_group_kernel<16, STAGE, PIPE>;
	default:
		return lz4e_decompress_group_kernel<8, STAGE, PIPE>;
	}
}

/* variant: LZ4E_DV_GROUP (synchronous copies) or LZ4E_DV_GROUP_PIPE */
static lz4e_dgroup_fn dgroup_fn(int group, int stage, int variant)
{
	const bool pipe = variant == LZ4E_DV_GROUP_PIPE;

	if (stage)
		return pipe ? dgroup_fn_gs<true, true>(group) : dgroup_fn_gs<true, false>(group);
	return pipe ? dgroup_fn_gs<false, true>(group) : dgroup_fn_gs<false, false>(group);
}

static int dgroup_norm(int group)
{
	return group == 4 || group == 16 ? group : 8;
}

static size_t dgroup_smem(int group, int stage, uint32_t stage_cap)
{
	const size_t per_group = stage ? (size_t)(stage_cap + LZ4E_D_PAD) : sizeof(lz4e_omap);

	return (size_t)LZ4E_D_WARPS * (32 / group) * per_group;
}

static int dgroup_per_sm(int group, int stage, uint32_t stage_cap, int variant)
{
	int per_sm = 0;
	lz4e_dgroup_fn fn = dgroup_fn(group, stage, variant);
	const size_t smem = dgroup_smem(group, stage, stage_cap);

	if (lz4e_set_smem((const void *)fn, smem))
		return -1;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, LZ4E_D_WARPS * 32, smem) !=
		    cudaSuccess ||
	    per_sm < 1)
		return -1;
	return per_sm;
}

static int launch_decompress_group(const lz4e_decompress_args *a, int group, int stage,
				   int variant, int sm_count, cudaStream_t stream,
				   int grid_limit, int *grid_out)
{
	group = dgroup_norm(group);
	const int per_sm = dgroup_per_sm(group, stage, a->stage_cap, variant);
	const int per_cta = LZ4E_D_WARPS * (32 / group);

	if (per_sm < 1)
		return -1;
	long long grid = (long long)per_sm * sm_count;
	const long long want = ((long long)a->nblocks + per_cta - 1) / per_cta;

	if (grid > want)
		grid = want;
	if (grid_limit > 0 && grid > grid_limit)
		grid = grid_limit;
	if (grid < 1)
		grid = 1;
	dgroup_fn(group, stage, variant)<<<(unsigned)grid, LZ4E_D_WARPS * 32,
					   dgroup_smem(group, stage, a->stage_cap), stream>>>(*a);
	if (grid_out)
		*grid_out = (int)grid;
	return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int lz4e_decompress_max_grid(int group, int stage, uint32_t stage_cap, int sm_count,
			     int variant)
{
	int per_sm = 0;

	if (variant == LZ4E_DV_WAVE) {
		per_sm = wave_per_sm();
		return per_sm < 1 ? -1 : per_sm * sm_count;
	}
	if (variant != LZ4E_DV_OLD && group != 1) {
		per_sm = dgroup_per_sm(dgroup_norm(group), stage, stage_cap, variant);
		return per_sm < 1 ? -1 : per_sm * sm_count;
	}

	if (group == 1) {
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
			    &per_sm, lz4e_decompress_lane_kernel, LZ4E_D_WARPS * 32, 0) !=
			    cudaSuccess ||
		    per_sm < 1)
			return -1;
		return per_sm * sm_count;
	}
	size_t smem = stage ? (size_t)LZ4E_D_WARPS * (32 / group) * (stage_cap + LZ4E_D_PAD) : 0;
	const void *fn;

	if (stage)
		fn = group == 8    ? (const void *)lz4e_decompress_kernel<8, true>
		     : group == 16 ? (const void *)lz4e_decompress_kernel<16, true>
				   : (const void *)lz4e_decompress_kernel<32, true>;
	else
		fn = group == 2	   ? (const void *)lz4e_decompress_kernel<2, false>
		     : group == 4  ? (const void *)lz4e_decompress_kernel<4, false>
		     : group == 8  ? (const void *)lz4e_decompress_kernel<8, false>
		     : group == 16 ? (const void *)lz4e_decompress_kernel<16, false>
				   : (const void *)lz4e_decompress_kernel<32, false>;
	if (lz4e_set_smem(fn, smem))
		return -1;
	if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, LZ4E_D_WARPS * 32,
							  smem) != cudaSuccess ||
	    per_sm < 1)
		return -1;
	return per_sm * sm_count;
}

int lz4e_launch_decompress(const lz4e_decompress_args *a, int group, int stage,
			   int variant, int sm_count, cudaStream_t stream, int grid_limit,
			   int *grid_out)
{
	if (variant == LZ4E_DV_WAVE)
		return launch_decompress_wave(a, sm_count, stream, grid_limit, grid_out);
	if (group == 1)
		return launch_decompress_lanes(a, sm_count, stream, grid_limit, grid_out);
	if (variant != LZ4E_DV_OLD)
		return launch_decompress_group(a, group, stage, variant, sm_count, stream,
					       grid_limit, grid_out);
	if (stage) {
		switch (group) {
		case 8:
			return launch_decompress_t<8, true>(a, sm_count, stream, grid_limit, grid_out);
		case 16:
			return launch_decompress_t<16, true>(a, sm_count, stream, grid_limit, grid_out);
		default:
			return launch_decompress_t<32, true>(a, sm_count, stream, grid_limit, grid_out);
		}
	}
	switch (group) {
	case 2:
		return launch_decompress_t<2, false>(a, sm_count, stream, grid_limit, grid_out);
	case 4:
		return launch_decompress_t<4, false>(a, sm_count, stream, grid_limit, grid_out);
	case 8:
		return launch_decompress_t<8, false>(a, sm_count, stream, grid_limit, grid_out);
	case 16:
		return launch_decompress_t<16, false>(a, sm_count, stream, grid_limit, grid_out);
	default:
		return launch_decompress_t<32, false>(a, sm_count, stream, grid_limit, grid_out);
	}
}
