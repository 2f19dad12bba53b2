/*
 * lz4e_simt.h — the few warp-level primitives the LZ4E kernels are written
 * against.
 *
 * Under nvcc these map 1:1 to the CUDA intrinsics.  The kernel BODIES
 * (lz4e_compress_body.cuh, lz4e_decompress_body.cuh) use only these names, so
 * tests/simt/ can also compile the very same source with g++ against a
 * coroutine-based 32-lane emulator and check the warp-cooperative logic
 * (speculative search batches, in-order commit, sub-warp groups) against the
 * oracle on machines without a GPU.  That emulator is TEST INFRASTRUCTURE: the
 * product library contains only the nvcc build and has no CPU path.
 */
#ifndef LZ4E_SIMT_H
#define LZ4E_SIMT_H

#include <stdint.h>

#if defined(__CUDACC__)

#define LZ4E_DEV __device__ __forceinline__
#define LZ4E_MEMBER __device__ __forceinline__ /* member functions */
/* rare paths are kept out of line so that the hot loops stay small enough for
 * the instruction cache */
#define LZ4E_DEV_NOINLINE __device__ __noinline__
#define LZ4E_FULL 0xffffffffu
/* address-space hint for pointers rebuilt from 64-bit descriptor fields */
#define LZ4E_ASSUME_GLOBAL(p) __builtin_assume(__isGlobal(p))
#define LZ4E_ASSUME_SHARED(p) __builtin_assume(__isShared(p))

LZ4E_DEV uint32_t lz4e_ballot(uint32_t mask, int pred)
{
	return __ballot_sync(mask, pred);
}
LZ4E_DEV uint32_t lz4e_shfl(uint32_t mask, uint32_t v, int src)
{
	return __shfl_sync(mask, v, src);
}
LZ4E_DEV uint32_t lz4e_match_any(uint32_t mask, uint32_t v)
{
	return __match_any_sync(mask, v);
}
LZ4E_DEV void lz4e_syncwarp(uint32_t mask)
{
	__syncwarp(mask);
}
LZ4E_DEV uint32_t lz4e_reduce_max(uint32_t mask, uint32_t v)
{
	return __reduce_max_sync(mask, v);
}
LZ4E_DEV uint32_t lz4e_funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)
{
	return __funnelshift_r(lo, hi, sh);
}
LZ4E_DEV int lz4e_clz(uint32_t v) { return __clz((int)v); }
LZ4E_DEV int lz4e_ffs(uint32_t v) { return __ffs((int)v); }
LZ4E_DEV int lz4e_popc(uint32_t v) { return __popc(v); }
LZ4E_DEV uint32_t lz4e_atomic_add(uint32_t *p, uint32_t v)
{
	return atomicAdd(p, v);
}
LZ4E_DEV void lz4e_atomic_or(uint32_t *p, uint32_t v) { atomicOr(p, v); }
LZ4E_DEV uint32_t lz4e_min(uint32_t a, uint32_t b) { return a < b ? a : b; }
LZ4E_DEV float lz4e_rcpf(float x) { return __frcp_rn(x); }

/*
 * 1-D TMA bulk copies global -> shared memory, completion counted on an
 * mbarrier in shared memory (cp.async.bulk, sm_90+; SASS: UBLKCP / SYNCS).
 * One thread issues, the copy engine of the SM moves the bytes, every thread
 * that wants to read them waits on the barrier's phase.
 */
#ifndef LZ4E_DISABLE_TMA
#define LZ4E_HAVE_TMA 1
#endif
LZ4E_DEV uint32_t lz4e_smem_addr(const void *p)
{
	return (uint32_t)__cvta_generic_to_shared(p);
}
LZ4E_DEV void lz4e_mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(lz4e_smem_addr(bar)), "r"(count)
		     : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
/* announce `bytes` of bulk-copy traffic for the current phase */
LZ4E_DEV void lz4e_mbar_expect(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(
			     lz4e_smem_addr(bar)),
		     "r"(bytes)
		     : "memory");
}
/* the issuing thread's arrival: the phase completes once the announced bytes
 * have landed */
LZ4E_DEV void lz4e_mbar_arrive(uint64_t *bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(lz4e_smem_addr(bar)) : "memory");
}
LZ4E_DEV void lz4e_mbar_wait(uint64_t *bar, uint32_t parity)
{
	asm volatile("{\n"
		     ".reg .pred p;\n"
		     "LZ4E_MBAR_WAIT: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		     "@p bra LZ4E_MBAR_DONE;\n"
		     "bra LZ4E_MBAR_WAIT;\n"
		     "LZ4E_MBAR_DONE:\n"
		     "}" ::"r"(lz4e_smem_addr(bar)),
		     "r"(parity)
		     : "memory");
}
/* earlier generic-proxy accesses to shared memory are ordered before later
 * async-proxy (TMA) writes */
LZ4E_DEV void lz4e_fence_async_smem(void)
{
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
/* dst (shared), src (global) 16-byte aligned, bytes a multiple of 16 */
LZ4E_DEV void lz4e_tma_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
			     lz4e_smem_addr(dst)),
		     "l"(src), "r"(bytes), "r"(lz4e_smem_addr(bar))
		     : "memory");
}

#else /* host build for tests/simt */

#include "simt_host.h"

#endif

#endif /* LZ4E_SIMT_H */
