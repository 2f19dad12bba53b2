/*
 * tests/simt/simt_host.h — TEST INFRASTRUCTURE, not product code.
 *
 * Host stand-ins for the warp primitives named in
 * lz4-sgori_b200/csrc/lz4e_simt.h, so the CUDA kernel bodies can be compiled
 * with g++ and executed by a 32-lane coroutine emulator (simt_run.cpp).  Lanes
 * are ucontext coroutines that only switch at collectives and are resumed in a
 * shuffled order, so a missing lz4e_syncwarp() between a store and another
 * lane's load shows up as a wrong result here as well.
 *
 * Nothing in the product library links this; it exists so that `-m "not gpu"`
 * tests can check the warp-cooperative algorithm against the oracle.
 */
#ifndef LZ4E_SIMT_HOST_H
#define LZ4E_SIMT_HOST_H

#include <stdint.h>

#define LZ4E_DEV static inline __attribute__((always_inline))
#define LZ4E_MEMBER inline
#define LZ4E_DEV_NOINLINE static __attribute__((noinline))
#define LZ4E_FULL 0xffffffffu
#define LZ4E_ASSUME_GLOBAL(p) ((void)(p))
#define LZ4E_ASSUME_SHARED(p) ((void)(p))

namespace simt
{
enum Op { OP_SYNC = 0, OP_BALLOT, OP_SHFL, OP_MATCH, OP_REDMAX };
uint32_t collective(Op op, uint32_t mask, uint32_t val, int src);
} // namespace simt

LZ4E_DEV uint32_t lz4e_ballot(uint32_t mask, int pred)
{
	return simt::collective(simt::OP_BALLOT, mask, pred ? 1u : 0u, 0);
}
LZ4E_DEV uint32_t lz4e_shfl(uint32_t mask, uint32_t v, int src)
{
	return simt::collective(simt::OP_SHFL, mask, v, src & 31);
}
LZ4E_DEV uint32_t lz4e_match_any(uint32_t mask, uint32_t v)
{
	return simt::collective(simt::OP_MATCH, mask, v, 0);
}
LZ4E_DEV uint32_t lz4e_reduce_max(uint32_t mask, uint32_t v)
{
	return simt::collective(simt::OP_REDMAX, mask, v, 0);
}
LZ4E_DEV void lz4e_syncwarp(uint32_t mask)
{
	(void)simt::collective(simt::OP_SYNC, mask, 0, 0);
}
LZ4E_DEV uint32_t lz4e_funnel_r(uint32_t lo, uint32_t hi, uint32_t sh)
{
	const uint64_t v = ((uint64_t)hi << 32) | lo;

	return (uint32_t)(v >> (sh & 31));
}
LZ4E_DEV int lz4e_clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
LZ4E_DEV int lz4e_ffs(uint32_t v) { return __builtin_ffs((int)v); }
LZ4E_DEV int lz4e_popc(uint32_t v) { return __builtin_popcount(v); }
LZ4E_DEV uint32_t lz4e_atomic_add(uint32_t *p, uint32_t v)
{
	return __atomic_fetch_add(p, v, __ATOMIC_RELAXED);
}
LZ4E_DEV void lz4e_atomic_or(uint32_t *p, uint32_t v)
{
	__atomic_fetch_or(p, v, __ATOMIC_RELAXED);
}
LZ4E_DEV uint32_t lz4e_min(uint32_t a, uint32_t b) { return a < b ? a : b; }
LZ4E_DEV float lz4e_rcpf(float x) { return 1.0f / x; }

#endif /* LZ4E_SIMT_HOST_H */
