/*
 * tests/simt/simt_run.cpp — TEST INFRASTRUCTURE, not product code.
 *
 * Runs the CUDA kernel bodies of lz4-sgori_b200/csrc/*_body.cuh on the CPU, one
 * emulated warp (32 ucontext coroutines) at a time.  "Device" addresses in the
 * descriptor tables are plain host addresses.  Used only by tests/.
 */
#include <ucontext.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#include "simt_host.h"

namespace simt
{
struct Warp {
	ucontext_t sched;
	ucontext_t ctx[32];
	std::vector<char> stack[32];
	bool finished[32];
	bool waiting[32];
	Op op[32];
	uint32_t mask[32];
	uint32_t val[32];
	int src[32];
	uint32_t result[32];
	int cur;
	uint32_t rng;
	std::function<void(uint32_t)> body;
};

static thread_local Warp *g_w;

static void die(const char *msg)
{
	fprintf(stderr, "simt emulator: %s\n", msg);
	abort();
}

uint32_t collective(Op op, uint32_t mask, uint32_t val, int src)
{
	Warp *w = g_w;
	const int l = w->cur;

	if (!((mask >> l) & 1u))
		die("lane called a collective with a mask that excludes it");
	w->op[l] = op;
	w->mask[l] = mask;
	w->val[l] = val;
	w->src[l] = src;
	w->waiting[l] = true;
	swapcontext(&w->ctx[l], &w->sched);
	return w->result[l];
}

static void trampoline()
{
	Warp *w = g_w;
	const int l = w->cur;

	w->body((uint32_t)l);
	w->finished[l] = true;
	swapcontext(&w->ctx[l], &w->sched);
}

static void resolve(Warp *w)
{
	for (int l = 0; l < 32; l++) {
		if (!w->waiting[l])
			continue;
		const uint32_t m = w->mask[l];
		bool ready = true;

		for (int j = 0; j < 32 && ready; j++) {
			if (!((m >> j) & 1u))
				continue;
			if (w->finished[j])
				die("collective names a lane that already exited");
			if (!w->waiting[j]) {
				ready = false;
			} else if (w->mask[j] != m || w->op[j] != w->op[l]) {
				/* the lane waits in another collective */
				ready = false;
			}
		}
		if (!ready)
			continue;
		uint32_t ballot = 0;

		for (int j = 0; j < 32; j++)
			if (((m >> j) & 1u) && w->val[j])
				ballot |= 1u << j;
		for (int j = 0; j < 32; j++) {
			if (!((m >> j) & 1u))
				continue;
			switch (w->op[j]) {
			case OP_SYNC:
				w->result[j] = 0;
				break;
			case OP_BALLOT:
				w->result[j] = ballot;
				break;
			case OP_SHFL:
				if (!((m >> w->src[j]) & 1u))
					die("shfl from a lane outside the mask");
				w->result[j] = w->val[w->src[j]];
				break;
			case OP_REDMAX: {
				uint32_t r = 0;

				for (int q = 0; q < 32; q++)
					if (((m >> q) & 1u) && w->val[q] > r)
						r = w->val[q];
				w->result[j] = r;
				break;
			}
			case OP_MATCH: {
				uint32_t r = 0;

				for (int q = 0; q < 32; q++)
					if (((m >> q) & 1u) && w->val[q] == w->val[j])
						r |= 1u << q;
				w->result[j] = r;
				break;
			}
			}
		}
		for (int j = 0; j < 32; j++)
			if ((m >> j) & 1u)
				w->waiting[j] = false;
	}
}

static void run_warp(const std::function<void(uint32_t)> &body, uint32_t seed)
{
	Warp *w = new Warp();

	w->body = body;
	w->rng = seed * 2654435761u + 12345u;
	g_w = w;
	for (int l = 0; l < 32; l++) {
		w->finished[l] = w->waiting[l] = false;
		w->stack[l].resize(256 * 1024);
		getcontext(&w->ctx[l]);
		w->ctx[l].uc_stack.ss_sp = w->stack[l].data();
		w->ctx[l].uc_stack.ss_size = w->stack[l].size();
		w->ctx[l].uc_link = &w->sched;
		makecontext(&w->ctx[l], trampoline, 0);
	}
	for (;;) {
		int order[32], nrun = 0;
		bool all_done = true;

		for (int l = 0; l < 32; l++) {
			if (!w->finished[l])
				all_done = false;
			if (!w->finished[l] && !w->waiting[l])
				order[nrun++] = l;
		}
		if (all_done)
			break;
		if (nrun == 0) {
			resolve(w);
			bool any = false;

			for (int l = 0; l < 32; l++)
				if (!w->finished[l] && !w->waiting[l])
					any = true;
			if (!any)
				die("deadlock: lanes wait in collectives that can never complete");
			continue;
		}
		for (int i = nrun - 1; i > 0; i--) { /* shuffled resume order */
			w->rng = w->rng * 1664525u + 1013904223u;
			const int j = (int)((w->rng >> 8) % (uint32_t)(i + 1));
			const int t = order[i];

			order[i] = order[j];
			order[j] = t;
		}
		for (int i = 0; i < nrun; i++) {
			w->cur = order[i];
			swapcontext(&w->sched, &w->ctx[order[i]]);
		}
		resolve(w);
	}
	g_w = nullptr;
	delete w;
}
} // namespace simt

#include "lz4e_compress_body.cuh"
#include "lz4e_decompress_body.cuh"
#include "lz4e_decompress_lane.cuh"
#include "lz4e_decompress_group.cuh"
#include "lz4e_decompress_wave.cuh"

extern "C" {

/* one emulated warp walks the whole batch (atomic work counter as on the GPU) */
int simt_compress_batch(const lz4e_b200_dev_bvec *src_bvecs,
			const lz4e_b200_dev_iter *src_iters,
			const lz4e_b200_dev_bvec *dst_bvecs,
			const lz4e_b200_dev_iter *dst_iters, uint32_t nblocks,
			uint32_t table_bytes, uint32_t in_cap, uint32_t out_cap,
			int32_t *rets, lz4e_b200_dev_end *ends, uint32_t *faults,
			uint32_t seed, uint32_t gin)
{
	lz4e_compress_args a;
	uint32_t counter = 0;
	const uint32_t bytes = lz4e_compress_smem_bytes(table_bytes, gin ? 0u : in_cap, out_cap);
	void *smraw = nullptr;
	/* GIN: the block is gathered into a global slot (stale garbage in it, as on
	 * the GPU when the slot is reused) */
	std::vector<uint8_t> slot(gin ? in_cap + 512 : 0, 0xA7);

	if (posix_memalign(&smraw, 1024, bytes + 64))
		return -1;
	memset(smraw, 0xCD, bytes + 64); /* stale garbage, as on the GPU */
	a.src_bvecs = src_bvecs;
	a.src_iters = src_iters;
	a.dst_bvecs = dst_bvecs;
	a.dst_iters = dst_iters;
	a.rets = rets;
	a.ends = ends;
	a.work_counter = &counter;
	a.order = nullptr;
	a.faults = faults;
	a.nblocks = nblocks;
	a.table_bytes = table_bytes;
	a.in_cap = in_cap;
	a.out_cap = out_cap;
	a.in_scratch = gin ? slot.data() : nullptr;
	a.in_stride = gin ? in_cap + 256 : 0;
	a.warps = 1;
	a.warp_smem = (bytes + 15u) & ~15u;
	a.tma = 0;
	uint8_t *sm = (uint8_t *)smraw;

	/* out_cap 64 = what lz4e_b200_compress_dev passes for linear destinations:
	 * output written in place, no staging */
	if (out_cap == 64 && gin)
		simt::run_warp([&](uint32_t lane) { lz4e_compress_warp<true, true>(a, sm, lane, 0); }, seed);
	else if (out_cap == 64)
		simt::run_warp([&](uint32_t lane) { lz4e_compress_warp<true, false>(a, sm, lane, 0); }, seed);
	else if (gin)
		simt::run_warp([&](uint32_t lane) { lz4e_compress_warp<false, true>(a, sm, lane, 0); }, seed);
	else
		simt::run_warp([&](uint32_t lane) { lz4e_compress_warp<false, false>(a, sm, lane, 0); }, seed);
	free(smraw);
	return 0;
}

int simt_decompress_batch(const lz4e_b200_dev_bvec *src_bvecs,
			  const lz4e_b200_dev_iter *src_iters,
			  const lz4e_b200_dev_bvec *dst_bvecs,
			  const lz4e_b200_dev_iter *dst_iters, uint32_t nblocks,
			  uint32_t group, uint32_t stage_cap, int32_t *rets,
			  uint32_t *faults, uint8_t *scratch, uint32_t scratch_stride,
			  uint32_t seed, uint32_t variant)
{
	lz4e_decompress_args a;
	uint32_t counter = 0;
	const uint32_t ngroups = 32 / (group ? group : 1);
	const uint32_t bytes = ngroups * (stage_cap + LZ4E_D_PAD) +
			       ngroups * (uint32_t)sizeof(lz4e_omap);
	void *smraw = nullptr;

	if (posix_memalign(&smraw, 1024, bytes + 64))
		return -1;
	memset(smraw, 0xCD, bytes + 64);
	a.src_bvecs = src_bvecs;
	a.src_iters = src_iters;
	a.dst_bvecs = dst_bvecs;
	a.dst_iters = dst_iters;
	a.rets = rets;
	a.work_counter = &counter;
	a.faults = faults;
	a.nblocks = nblocks;
	a.stage_cap = stage_cap;
	a.scratch = scratch;
	a.scratch_stride = scratch_stride;
	a.in_scratch = nullptr;
	a.in_stride = 0;
	a.order = nullptr;
	a.wave_serial = 14;
	uint8_t *sm = (uint8_t *)smraw;

	if (group == 1) {
		/* lane-per-block decoder: `scratch` holds 32 output slots, the
		 * input slots are allocated here */
		lz4e_lane_smem *ls = (lz4e_lane_smem *)calloc(1, sizeof(*ls));
		uint8_t *ins = (uint8_t *)malloc((size_t)32 * (scratch_stride + 64) + 64);

		a.in_scratch = ins;
		a.in_stride = scratch_stride + 64;
		simt::run_warp([&](uint32_t lane) { lz4e_decompress_lanes(a, ls, 0, lane); }, seed);
		free(ins);
		free(ls);
		free(smraw);
		return 0;
	}

	if (variant == 3) {
		/* lz4e_decompress_wave.cuh: one warp per block; `scratch` holds one
		 * output slot */
		lz4e_wave_smem *wsm = nullptr;

		if (posix_memalign((void **)&wsm, 64, sizeof(*wsm)))
			return -1;
		memset(wsm, 0xCD, sizeof(*wsm));
		simt::run_warp([&](uint32_t lane) { lz4e_decompress_wave(a, wsm, 0, lane); }, seed);
		free(wsm);
		free(smraw);
		return 0;
	}

	if (variant) {
		/* lz4e_decompress_group.cuh: variant 1 synchronous, 2 pipelined */
#define RUN_G(GS)                                                                        \
	simt::run_warp(                                                                  \
		[&](uint32_t lane) {                                                     \
			if (stage_cap && variant == 2)                                   \
				lz4e_decompress_groups<GS, true, true>(a, sm, 0, lane);  \
			else if (stage_cap)                                              \
				lz4e_decompress_groups<GS, true, false>(a, sm, 0, lane); \
			else if (variant == 2)                                           \
				lz4e_decompress_groups<GS, false, true>(a, sm, 0, lane); \
			else                                                             \
				lz4e_decompress_groups<GS, false, false>(a, sm, 0, lane); \
		},                                                                       \
		seed)
		if (group == 4)
			RUN_G(4);
		else if (group == 16)
			RUN_G(16);
		else
			RUN_G(8);
#undef RUN_G
		free(smraw);
		return 0;
	}

	simt::run_warp(
		[&](uint32_t lane) {
			switch (group) {
			case 2:
				lz4e_decompress_warp<2>(a, sm, 0, lane);
				break;
			case 4:
				lz4e_decompress_warp<4>(a, sm, 0, lane);
				break;
			case 8:
				lz4e_decompress_warp<8>(a, sm, 0, lane);
				break;
			case 16:
				lz4e_decompress_warp<16>(a, sm, 0, lane);
				break;
			default:
				lz4e_decompress_warp<32>(a, sm, 0, lane);
				break;
			}
		},
		seed);
	free(smraw);
	return 0;
}
}
