/*
 * lz4e_abi.cu — implementation of the C ABI declared in include/lz4e_b200.h:
 * context, page arenas, raw memory helpers and the device-level batch entry
 * points.  Everything here either enqueues CUDA work or fails with an error
 * code; there is no CPU implementation of the codec in this library.
 */
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/lz4e_b200.h"
#include "lz4e_compress_body.cuh"
#include "lz4e_decompress_body.cuh"
#include "lz4e_kernels.h"

#define CTL_COUNTERS 256
#define CTL_FAULTS CTL_COUNTERS
#define CTL_WORDS (CTL_COUNTERS + 8)

/* held while a *_dev call plans and enqueues its launch: kernel attributes
 * (dynamic shared memory size) are per function and device, not per launch */
static std::mutex g_launch_mu;

struct lz4e_b200_arena {
	lz4e_b200_ctx *ctx;
	uint8_t *host;
	uint8_t *dev;
	uint8_t *host_raw, *dev_raw; /* as allocated; host/dev are page aligned */
	size_t npages;
};

struct lz4e_b200_ctx {
	int device;
	int sm_count;
	cudaStream_t stream;
	uint32_t *d_ctl; /* work counters + fault word */
	std::atomic<uint32_t> ctl_next;
	std::atomic<uint64_t> launches;
	/* work buffers, one per (stream, purpose) in use: launches on different
	 * streams never share slots */
	struct cslot {
		cudaStream_t stream;
		int kind; /* SCRATCH_* */
		uint8_t *p;
		size_t bytes;
	};
	std::vector<cslot> cscratch;
	int gin_min;	 /* blocks above this many bytes use the GIN compress kernels */
	int gin_per_sm;	 /* resident warps per SM for them (0: as many as fit) */
	int gin_small;	 /* batches of up to this many waves of resident shared-memory
			  * warps keep their input in shared memory (0: never) */
	std::mutex mu;
	std::vector<lz4e_b200_arena *> arenas;
	char err[256];
	int dgroup;  /* decompression group size: 1 (lane per block), 4, 8, 16; 0 auto */
	int dstage;  /* -1 auto, 0 never stage, 1 stage when it fits */
	int dvariant; /* LZ4E_DV_*: which decoder; -1 auto */
	int dorder;   /* cost-ordered work queue from this many blocks on; 0 = never */
	int wave_min; /* blocks above 8 KiB: wave decoder from this many blocks on */
	int lane_per_sm; /* lane-per-block decoder: CTAs per SM (0: as many as fit) */
	int wave_serial; /* wave decoder: sequence-serial copy from this many bytes per sequence */
	int p12;      /* compress: packed 12-bit table for blocks <= 4 KiB */
	int tma;      /* compress: TMA bulk staging of the input (shared-memory kernels) */
	/* bytes moved between host and device through this context: copy engine
	 * (host counters) and lz4e_xfer_kernel (device counters d_moved[0..1]) */
	std::atomic<uint64_t> h2d_bytes, d2h_bytes;
	unsigned long long *d_moved;
	/* state other layers hang on the context (host workspace, chunk pool):
	 * created once, torn down by lz4e_b200_ctx_destroy */
	struct attachment {
		int slot;
		void *p;
		void (*dtor)(lz4e_b200_ctx *, void *);
	};
	std::vector<attachment> attached;
};

static void set_err(lz4e_b200_ctx *ctx, const char *fmt, ...)
{
	va_list ap;

	if (!ctx)
		return;
	va_start(ap, fmt);
	vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
	va_end(ap);
}

#define CU(ctx, call)                                                        \
	do {                                                                 \
		cudaError_t e_ = (call);                                     \
		if (e_ != cudaSuccess) {                                     \
			set_err(ctx, "%s failed: %s", #call,                 \
				cudaGetErrorString(e_));                     \
			return LZ4E_B200_EIO;                                \
		}                                                            \
	} while (0)

static cudaStream_t pick(lz4e_b200_ctx *ctx, void *stream)
{
	return stream ? (cudaStream_t)stream : ctx->stream;
}

extern "C" {

int lz4e_b200_device_count(void)
{
	int n = 0;

	if (cudaGetDeviceCount(&n) != cudaSuccess)
		return 0;
	return n;
}

int lz4e_b200_ctx_create(int device, lz4e_b200_ctx **out)
{
	cudaDeviceProp prop;
	int n = 0;

	if (!out)
		return LZ4E_B200_EINVAL;
	*out = nullptr;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
		fprintf(stderr,
			"lz4e_b200: no usable CUDA device %d (found %d); this library "
			"has no CPU path\n",
			device, n);
		return LZ4E_B200_ENODEV;
	}
	if (cudaSetDevice(device) != cudaSuccess ||
	    cudaGetDeviceProperties(&prop, device) != cudaSuccess)
		return LZ4E_B200_ENODEV;
	if (prop.major < 10) {
		fprintf(stderr,
			"lz4e_b200: device %d is sm_%d%d; kernels are built for sm_100a only\n",
			device, prop.major, prop.minor);
		return LZ4E_B200_ENODEV;
	}
	lz4e_b200_ctx *ctx = new lz4e_b200_ctx();

	ctx->device = device;
	ctx->sm_count = prop.multiProcessorCount;
	ctx->ctl_next = 0;
	ctx->launches = 0;
	ctx->err[0] = 0;
	ctx->dgroup = 0; /* auto: by block size */
	ctx->dstage = -1;
	if (const char *e = getenv("LZ4E_B200_DGROUP")) {
		int g = atoi(e);

		if (g == 1 || g == 2 || g == 4 || g == 8 || g == 16 || g == 32)
			ctx->dgroup = g;
	}
	if (const char *e = getenv("LZ4E_B200_DSTAGE"))
		ctx->dstage = atoi(e) ? 1 : 0;
	ctx->dvariant = -1;
	if (const char *e = getenv("LZ4E_B200_DVARIANT"))
		ctx->dvariant = atoi(e);
	ctx->wave_serial = 32;
	if (const char *e = getenv("LZ4E_B200_WAVE_SERIAL"))
		ctx->wave_serial = atoi(e);
	ctx->p12 = 0;
	if (const char *e = getenv("LZ4E_B200_P12"))
		ctx->p12 = atoi(e) ? 1 : 0;
	ctx->tma = 0;
	if (const char *e = getenv("LZ4E_B200_TMA"))
		ctx->tma = atoi(e) ? 1 : 0;
	ctx->wave_min = 0;
	if (const char *e = getenv("LZ4E_B200_WAVE_MIN"))
		ctx->wave_min = atoi(e);
	ctx->lane_per_sm = 0;
	if (const char *e = getenv("LZ4E_B200_LANE_PER_SM"))
		ctx->lane_per_sm = atoi(e);
	ctx->dorder = 2048;
	if (const char *e = getenv("LZ4E_B200_DORDER"))
		ctx->dorder = atoi(e);
	ctx->gin_min = 12288;
	ctx->gin_per_sm = 0;
	ctx->gin_small = 2;
	if (const char *e = getenv("LZ4E_B200_GIN_SMALL"))
		ctx->gin_small = atoi(e);
	if (const char *e = getenv("LZ4E_B200_GIN_MIN"))
		ctx->gin_min = atoi(e);
	if (const char *e = getenv("LZ4E_B200_GIN_PER_SM"))
		ctx->gin_per_sm = atoi(e);
	ctx->h2d_bytes = 0;
	ctx->d2h_bytes = 0;
	ctx->d_moved = nullptr;
	if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaMalloc((void **)&ctx->d_ctl, CTL_WORDS * sizeof(uint32_t)) != cudaSuccess ||
	    cudaMemset(ctx->d_ctl, 0, CTL_WORDS * sizeof(uint32_t)) != cudaSuccess ||
	    cudaMalloc((void **)&ctx->d_moved, 2 * sizeof(unsigned long long)) != cudaSuccess ||
	    cudaMemset(ctx->d_moved, 0, 2 * sizeof(unsigned long long)) != cudaSuccess) {
		fprintf(stderr, "lz4e_b200: context setup failed: %s\n",
			cudaGetErrorString(cudaGetLastError()));
		delete ctx;
		return LZ4E_B200_EIO;
	}
	*out = ctx;
	return LZ4E_B200_OK;
}

void lz4e_b200_ctx_destroy(lz4e_b200_ctx *ctx)
{
	if (!ctx)
		return;
	cudaSetDevice(ctx->device);
	cudaDeviceSynchronize();
	/* layers above first (they may still own arenas and streams), newest first */
	while (!ctx->attached.empty()) {
		lz4e_b200_ctx::attachment at = ctx->attached.back();

		ctx->attached.pop_back();
		if (at.dtor)
			at.dtor(ctx, at.p);
		free(at.p);
	}
	/* arenas the caller left behind go with the context: their handles are
	 * invalid from here on (documented in lz4e_b200.h) */
	while (!ctx->arenas.empty())
		lz4e_b200_arena_destroy(ctx->arenas.back());
	for (auto &c : ctx->cscratch)
		cudaFree(c.p);
	cudaFree(ctx->d_ctl);
	cudaFree(ctx->d_moved);
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}

lz4e_b200_ctx *lz4e_b200_default_ctx(void)
{
	static std::mutex mu;
	static lz4e_b200_ctx *def;
	std::lock_guard<std::mutex> lk(mu);

	if (!def) {
		const char *e = getenv("LZ4E_B200_DEVICE");

		if (lz4e_b200_ctx_create(e ? atoi(e) : 0, &def) != LZ4E_B200_OK)
			def = nullptr;
	}
	return def;
}

const char *lz4e_b200_strerror(lz4e_b200_ctx *ctx)
{
	return ctx ? ctx->err : "no context";
}

int lz4e_b200_ctx_tune(lz4e_b200_ctx *ctx, const char *key, int value)
{
	if (!ctx || !key)
		return LZ4E_B200_EINVAL;
	if (!strcmp(key, "dgroup"))
		ctx->dgroup = value;
	else if (!strcmp(key, "dstage"))
		ctx->dstage = value;
	else if (!strcmp(key, "dvariant"))
		ctx->dvariant = value;
	else if (!strcmp(key, "dorder"))
		ctx->dorder = value;
	else if (!strcmp(key, "wave_min"))
		ctx->wave_min = value;
	else if (!strcmp(key, "lane_per_sm"))
		ctx->lane_per_sm = value;
	else if (!strcmp(key, "wave_serial"))
		ctx->wave_serial = value;
	else if (!strcmp(key, "p12"))
		ctx->p12 = value ? 1 : 0;
	else if (!strcmp(key, "tma"))
		ctx->tma = value ? 1 : 0;
	else if (!strcmp(key, "gin_min"))
		ctx->gin_min = value;
	else if (!strcmp(key, "gin_per_sm"))
		ctx->gin_per_sm = value;
	else if (!strcmp(key, "gin_small"))
		ctx->gin_small = value;
	else
		return LZ4E_B200_EINVAL;
	return LZ4E_B200_OK;
}

int lz4e_b200_ctx_device(lz4e_b200_ctx *ctx) { return ctx->device; }
int lz4e_b200_ctx_sm_count(lz4e_b200_ctx *ctx) { return ctx->sm_count; }
uint64_t lz4e_b200_kernel_launches(lz4e_b200_ctx *ctx) { return ctx->launches.load(); }

int lz4e_b200_sync(lz4e_b200_ctx *ctx, void *stream)
{
	uint32_t faults = 0;

	cudaSetDevice(ctx->device);
	CU(ctx, cudaStreamSynchronize(pick(ctx, stream)));
	CU(ctx, cudaMemcpy(&faults, ctx->d_ctl + CTL_FAULTS, 4, cudaMemcpyDeviceToHost));
	if (faults) {
		CU(ctx, cudaMemset(ctx->d_ctl + CTL_FAULTS, 0, 4));
		set_err(ctx, "device fault flags 0x%x: a block exceeded the limits or "
			     "promises given to a *_dev call",
			faults);
		return LZ4E_B200_ERANGE;
	}
	return LZ4E_B200_OK;
}

/* ---- raw memory ---------------------------------------------------------- */
void *lz4e_b200_dev_alloc(lz4e_b200_ctx *ctx, size_t bytes)
{
	void *p = nullptr;

	cudaSetDevice(ctx->device);
	if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
		set_err(ctx, "cudaMalloc(%zu) failed", bytes);
		return nullptr;
	}
	return p;
}

void lz4e_b200_dev_free(lz4e_b200_ctx *ctx, void *p)
{
	cudaSetDevice(ctx->device);
	cudaFree(p);
}

void *lz4e_b200_host_alloc(lz4e_b200_ctx *ctx, size_t bytes)
{
	void *p = nullptr;

	cudaSetDevice(ctx->device);
	if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
		set_err(ctx, "cudaHostAlloc(%zu) failed", bytes);
		return nullptr;
	}
	return p;
}

void lz4e_b200_host_free(lz4e_b200_ctx *ctx, void *p)
{
	cudaSetDevice(ctx->device);
	cudaFreeHost(p);
}

int lz4e_b200_copy_h2d(lz4e_b200_ctx *ctx, void *stream, void *dst_dev,
		       const void *src_host, size_t bytes)
{
	cudaSetDevice(ctx->device);
	CU(ctx, cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice,
				pick(ctx, stream)));
	ctx->h2d_bytes += bytes;
	return LZ4E_B200_OK;
}

int lz4e_b200_copy_d2h(lz4e_b200_ctx *ctx, void *stream, void *dst_host,
		       const void *src_dev, size_t bytes)
{
	cudaSetDevice(ctx->device);
	CU(ctx, cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost,
				pick(ctx, stream)));
	ctx->d2h_bytes += bytes;
	return LZ4E_B200_OK;
}

/* ---- streams and events --------------------------------------------------- */
void *lz4e_b200_stream_create(lz4e_b200_ctx *ctx)
{
	cudaStream_t s = nullptr;

	cudaSetDevice(ctx->device);
	if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
		set_err(ctx, "cudaStreamCreate failed");
		return nullptr;
	}
	return (void *)s;
}

void lz4e_b200_stream_destroy(lz4e_b200_ctx *ctx, void *stream)
{
	cudaSetDevice(ctx->device);
	if (stream)
		cudaStreamDestroy((cudaStream_t)stream);
}

void *lz4e_b200_event_create(lz4e_b200_ctx *ctx)
{
	cudaEvent_t e = nullptr;

	cudaSetDevice(ctx->device);
	if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
		set_err(ctx, "cudaEventCreate failed");
		return nullptr;
	}
	return (void *)e;
}

void *lz4e_b200_event_create_timed(lz4e_b200_ctx *ctx)
{
	cudaEvent_t e = nullptr;

	cudaSetDevice(ctx->device);
	if (cudaEventCreate(&e) != cudaSuccess) {
		set_err(ctx, "cudaEventCreate failed");
		return nullptr;
	}
	return (void *)e;
}

/* milliseconds from `from` to `to` (both recorded and completed), < 0 on error */
float lz4e_b200_event_elapsed_ms(lz4e_b200_ctx *ctx, void *from, void *to)
{
	float ms = -1.0f;

	cudaSetDevice(ctx->device);
	if (cudaEventElapsedTime(&ms, (cudaEvent_t)from, (cudaEvent_t)to) != cudaSuccess)
		return -1.0f;
	return ms;
}

void lz4e_b200_event_destroy(lz4e_b200_ctx *ctx, void *event)
{
	cudaSetDevice(ctx->device);
	if (event)
		cudaEventDestroy((cudaEvent_t)event);
}

int lz4e_b200_event_record(lz4e_b200_ctx *ctx, void *event, void *stream)
{
	cudaSetDevice(ctx->device);
	CU(ctx, cudaEventRecord((cudaEvent_t)event, pick(ctx, stream)));
	return LZ4E_B200_OK;
}

int lz4e_b200_stream_wait_event(lz4e_b200_ctx *ctx, void *stream, void *event)
{
	cudaSetDevice(ctx->device);
	CU(ctx, cudaStreamWaitEvent(pick(ctx, stream), (cudaEvent_t)event, 0));
	return LZ4E_B200_OK;
}

/* ---- arenas -------------------------------------------------------------- */
int lz4e_b200_arena_create(lz4e_b200_ctx *ctx, size_t npages, lz4e_b200_arena **out)
{
	lz4e_b200_arena *a;

	if (!ctx || !out || !npages)
		return LZ4E_B200_EINVAL;
	cudaSetDevice(ctx->device);
	a = new lz4e_b200_arena();
	a->ctx = ctx;
	a->npages = npages;
	a->host = a->dev = a->host_raw = a->dev_raw = nullptr;
	/* neither allocator promises page alignment for small sizes */
	if (cudaHostAlloc((void **)&a->host_raw, (npages + 1) * PAGE_SIZE,
			  cudaHostAllocDefault) != cudaSuccess ||
	    cudaMalloc((void **)&a->dev_raw, (npages + 1) * PAGE_SIZE) != cudaSuccess) {
		set_err(ctx, "arena of %zu pages: %s", npages,
			cudaGetErrorString(cudaGetLastError()));
		cudaFreeHost(a->host_raw);
		cudaFree(a->dev_raw);
		delete a;
		return LZ4E_B200_ENOMEM;
	}
	a->host = (uint8_t *)(((uintptr_t)a->host_raw + PAGE_SIZE - 1) &
			      ~(uintptr_t)(PAGE_SIZE - 1));
	a->dev = (uint8_t *)(((uintptr_t)a->dev_raw + PAGE_SIZE - 1) &
			     ~(uintptr_t)(PAGE_SIZE - 1));
	{
		std::lock_guard<std::mutex> lk(ctx->mu);
		ctx->arenas.push_back(a);
	}
	*out = a;
	return LZ4E_B200_OK;
}

void lz4e_b200_arena_destroy(lz4e_b200_arena *a)
{
	if (!a)
		return;
	lz4e_b200_ctx *ctx = a->ctx;
	{
		std::lock_guard<std::mutex> lk(ctx->mu);
		for (size_t i = 0; i < ctx->arenas.size(); i++)
			if (ctx->arenas[i] == a) {
				ctx->arenas.erase(ctx->arenas.begin() + i);
				break;
			}
	}
	cudaSetDevice(ctx->device);
	cudaFreeHost(a->host_raw);
	cudaFree(a->dev_raw);
	delete a;
}

struct page *lz4e_b200_arena_pages(lz4e_b200_arena *a) { return (struct page *)a->host; }
void *lz4e_b200_arena_dev_base(lz4e_b200_arena *a) { return a->dev; }
size_t lz4e_b200_arena_npages(lz4e_b200_arena *a) { return a->npages; }

int lz4e_b200_arena_upload(lz4e_b200_arena *a, size_t first_page, size_t npages)
{
	if (first_page + npages > a->npages)
		return LZ4E_B200_EINVAL;
	return lz4e_b200_copy_h2d(a->ctx, nullptr, a->dev + first_page * PAGE_SIZE,
				  a->host + first_page * PAGE_SIZE, npages * PAGE_SIZE);
}

int lz4e_b200_arena_download(lz4e_b200_arena *a, size_t first_page, size_t npages)
{
	if (first_page + npages > a->npages)
		return LZ4E_B200_EINVAL;
	return lz4e_b200_copy_d2h(a->ctx, nullptr, a->host + first_page * PAGE_SIZE,
				  a->dev + first_page * PAGE_SIZE, npages * PAGE_SIZE);
}

size_t lz4e_b200_arena_snapshot(lz4e_b200_ctx *ctx, lz4e_b200_arena **out, size_t max)
{
	std::lock_guard<std::mutex> lk(ctx->mu);
	size_t n = 0;

	for (lz4e_b200_arena *a : ctx->arenas) {
		if (n == max)
			break;
		out[n++] = a;
	}
	return n;
}

uint64_t lz4e_b200_arena_translate(lz4e_b200_ctx *ctx, const void *p,
				   lz4e_b200_arena **arena_out)
{
	std::lock_guard<std::mutex> lk(ctx->mu);

	for (lz4e_b200_arena *a : ctx->arenas) {
		const uint8_t *q = (const uint8_t *)p;

		if (q >= a->host && q < a->host + a->npages * PAGE_SIZE) {
			if (arena_out)
				*arena_out = a;
			return (uint64_t)(uintptr_t)(a->dev + (q - a->host));
		}
	}
	return 0;
}

/* ---- device-level batches ------------------------------------------------ */
static uint32_t *next_counter(lz4e_b200_ctx *ctx, cudaStream_t s)
{
	uint32_t *c = ctx->d_ctl + (ctx->ctl_next.fetch_add(1) % CTL_COUNTERS);

	cudaMemsetAsync(c, 0, sizeof(uint32_t), s);
	return c;
}

enum { SCRATCH_COMPRESS_IN = 0, SCRATCH_DECOMPRESS = 1, SCRATCH_COMPACT = 2, SCRATCH_ORDER = 3 };

/*
 * Work buffer of `bytes` bytes for a launch of the given purpose on stream s.
 * Buffers are kept per (stream, purpose): two launches on different streams
 * never share slots, and work already enqueued on s keeps its buffer (growing
 * waits for the stream first).
 */
static uint8_t *stream_scratch(lz4e_b200_ctx *ctx, cudaStream_t s, int kind, size_t bytes)
{
	std::lock_guard<std::mutex> lk(ctx->mu);

	for (auto &c : ctx->cscratch) {
		if (c.stream != s || c.kind != kind)
			continue;
		if (c.bytes >= bytes)
			return c.p;
		if (cudaStreamSynchronize(s) != cudaSuccess)
			return nullptr;
		cudaFree(c.p);
		c.p = nullptr;
		c.bytes = 0;
		if (cudaMalloc((void **)&c.p, bytes) != cudaSuccess) {
			c.p = nullptr;
			return nullptr;
		}
		c.bytes = bytes;
		return c.p;
	}
	lz4e_b200_ctx::cslot c = { s, kind, nullptr, 0 };

	if (cudaMalloc((void **)&c.p, bytes) != cudaSuccess)
		return nullptr;
	c.bytes = bytes;
	ctx->cscratch.push_back(c);
	return c.p;
}

/* internal helpers for the other translation units */
uint8_t *lz4e_ctx_scratch(lz4e_b200_ctx *ctx, void *stream, size_t bytes)
{
	return stream_scratch(ctx, pick(ctx, stream), SCRATCH_COMPACT, bytes);
}
uint32_t *lz4e_ctx_faults(lz4e_b200_ctx *ctx) { return ctx->d_ctl + CTL_FAULTS; }
void *lz4e_ctx_stream(lz4e_b200_ctx *ctx, void *stream) { return (void *)pick(ctx, stream); }
void lz4e_ctx_count_launches(lz4e_b200_ctx *ctx, unsigned k) { ctx->launches += k; }

int lz4e_b200_compress_dev(lz4e_b200_ctx *ctx, void *stream,
			   const struct lz4e_b200_dev_bvec *src_bvecs,
			   const struct lz4e_b200_dev_iter *src_iters,
			   const struct lz4e_b200_dev_bvec *dst_bvecs,
			   const struct lz4e_b200_dev_iter *dst_iters,
			   uint32_t nblocks, uint32_t max_in_bytes, uint32_t flags,
			   int32_t *rets, struct lz4e_b200_dev_end *ends)
{
	lz4e_compress_args a;
	cudaStream_t s;

	if (!ctx)
		return LZ4E_B200_ENODEV;
	if (nblocks == 0)
		return LZ4E_B200_OK;
	if (max_in_bytes > LZ4E_B200_MAX_BLOCK) {
		set_err(ctx, "blocks above %u bytes (%u) are not supported by the "
			     "compressor", (unsigned)LZ4E_B200_MAX_BLOCK, max_in_bytes);
		return LZ4E_B200_EINVAL;
	}
	cudaSetDevice(ctx->device);
	std::lock_guard<std::mutex> launch_lock(g_launch_mu);
	s = pick(ctx, stream);
	a.src_bvecs = src_bvecs;
	a.src_iters = src_iters;
	a.dst_bvecs = dst_bvecs;
	a.dst_iters = dst_iters;
	a.rets = rets;
	a.ends = ends;
	a.work_counter = next_counter(ctx, s);
	a.faults = ctx->d_ctl + CTL_FAULTS;
	a.nblocks = nblocks;
	a.tma = (uint32_t)ctx->tma;
	a.table_bytes = (flags & LZ4E_B200_DEV_NO_BYU16) ? 8192u : 16384u;
	/* blocks of at most 4 KiB: 12-bit positions, 8192 entries in 12 KiB
	 * (lz4e_table<.., P12>): 14 instead of 11 resident warps per SM.  An
	 * option (LZ4E_B200_P12=1): sparse pages gain 13 %, but text loses 16 %
	 * to the longer table update on its serial chain and the mixed batch
	 * ends at 93.5 instead of 101.6 GB/s (profiles/r2ai_packed_table.md) */
	if (!(flags & LZ4E_B200_DEV_NO_BYU16) && max_in_bytes <= 4096u && ctx->p12)
		a.table_bytes = LZ4E_K_P12_BYTES;
	if (max_in_bytes > 65536u) /* 32-bit positions: 4096 entries of 4 bytes */
		a.table_bytes = 16384u;
	a.in_cap = (max_in_bytes + 15u) & ~15u;
	if (a.in_cap < 16)
		a.in_cap = 16;
	{
		/* whole output staged when it is small (one coalesced flush per
		 * block), a 2 KiB streaming window otherwise */
		const uint32_t bound = max_in_bytes + max_in_bytes / 255u + 16u;

		a.out_cap = bound <= 4352u ? ((bound + 15u) & ~15u) + 16u : 1024u;
		if (a.out_cap < 1024u)
			a.out_cap = 1024u;
		/* the window holds any length run by itself: one length byte per
		 * 255 bytes of a literal run or match, 4 KiB for 1 MiB blocks,
		 * 64 KiB for 16 MiB ones (blocks up to 255 KiB: the 1 KiB above) */
		if (a.out_cap < ((max_in_bytes / 255u + 20u + 15u) & ~15u))
			a.out_cap = (max_in_bytes / 255u + 20u + 15u) & ~15u;
		/* linear destinations are written in place: no staging, more
		 * resident warps (4 KiB blocks: 9 -> 11 per SM) */
		if (flags & LZ4E_B200_DEV_DST_LINEAR)
			a.out_cap = 64u;
	}
	{
		/* large blocks: input in a global slot read through L1 instead of
		 * shared memory (64 KiB blocks: 3 -> 16 resident warps per SM) */
		const int direct = (flags & LZ4E_B200_DEV_DST_LINEAR) ? 1 : 0;
		int gin = (int)max_in_bytes > ctx->gin_min;

		/* a batch small enough to be resident all at once with its input
		 * in SHARED memory runs there: with few warps per SM occupancy is
		 * no concern and shared-memory loads beat loads through L1/L2 */
		{
			const size_t per_warp =
				lz4e_compress_smem_bytes(a.table_bytes, a.in_cap, a.out_cap);
			const size_t cta_max = (size_t)227 * 1024 - 1024;

			if (per_warp > cta_max)
				gin = 1;
			else if (gin && ctx->gin_small > 0 &&
				 (size_t)nblocks <= (size_t)ctx->sm_count * (cta_max / per_warp) *
							    (size_t)ctx->gin_small)
				gin = 0;
		}
		int grid = 0;
		size_t smem = 0;

		a.in_scratch = nullptr;
		a.in_stride = 0;
		if (lz4e_compress_plan(&a, direct, gin, ctx->sm_count, gin ? ctx->gin_per_sm : 0,
				       &grid, &smem)) {
			set_err(ctx, "compress launch plan failed: %s",
				cudaGetErrorString(cudaGetLastError()));
			return LZ4E_B200_EIO;
		}
		if (gin) {
			/* input slots of one launch stay below 4 GiB */
			const size_t budget = (size_t)4 << 30;

			a.in_stride = ((a.in_cap + 63u) & ~255u) + 256u;
			if ((size_t)grid * a.warps * a.in_stride > budget) {
				grid = (int)(budget / a.in_stride / a.warps);
				if (grid < 1)
					grid = 1;
			}
			a.in_scratch = stream_scratch(ctx, s, SCRATCH_COMPRESS_IN,
						      (size_t)grid * a.warps * a.in_stride + 256);
			if (!a.in_scratch) {
				set_err(ctx, "compress input slots (%zu bytes): out of device memory",
					(size_t)grid * a.in_stride);
				return LZ4E_B200_ENOMEM;
			}
		}
		/* bios of very different sizes (configs[4]: 4-128 KiB): hand the
		 * large ones out first, or the batch ends with one warp working on
		 * a 128 KiB block it took last while the machine idles */
		a.order = nullptr;
		if (max_in_bytes > 8192u && nblocks >= 32u && ctx->dorder > 0) {
			uint32_t *ord = (uint32_t *)stream_scratch(
				ctx, s, SCRATCH_ORDER,
				((size_t)nblocks + 2 * LZ4E_ORDER_BUCKETS) * sizeof(uint32_t));

			if (ord && !lz4e_launch_order(src_iters, nullptr, nblocks, ord + nblocks, ord, s,
						       max_in_bytes + 1u)) {
				a.order = ord;
				ctx->launches += 2;
			}
		}
		if (lz4e_launch_compress(&a, direct, gin, grid, smem, s)) {
			set_err(ctx, "compress launch failed: %s",
				cudaGetErrorString(cudaGetLastError()));
			return LZ4E_B200_EIO;
		}
	}
	ctx->launches++;
	return LZ4E_B200_OK;
}

int lz4e_b200_decompress_dev(lz4e_b200_ctx *ctx, void *stream,
			     const struct lz4e_b200_dev_bvec *src_bvecs,
			     const struct lz4e_b200_dev_iter *src_iters,
			     const struct lz4e_b200_dev_bvec *dst_bvecs,
			     const struct lz4e_b200_dev_iter *dst_iters,
			     uint32_t nblocks, uint32_t max_in_bytes,
			     uint32_t max_out_bytes, uint32_t flags, int32_t *rets)
{
	lz4e_decompress_args a;
	cudaStream_t s;
	int stage, grid_limit = 0, group, variant;
	const bool dst_linear = flags & LZ4E_B200_DEV_DST_LINEAR;
	const bool small = max_out_bytes <= 8192u;

	if (!ctx)
		return LZ4E_B200_ENODEV;
	if (nblocks == 0)
		return LZ4E_B200_OK;
	/*
	 * Which decoder (measured on B200, profiles/r2d_decoder_variants.md):
	 *  - blocks above 8 KiB: the group decoder of lz4e_decompress_group.cuh
	 *    with pipelined match heads, 16 lanes per block while the batch
	 *    cannot fill the resident groups of 8 (the serial chain of one
	 *    block bounds the launch), 8 lanes otherwise;
	 *  - small blocks, batch large enough to give every resident LANE a
	 *    block: the lane-per-block decoder;
	 *  - small blocks otherwise: the group decoder, 8 lanes, output staged in
	 *    shared memory when the destination is scattered or the batch fits
	 *    one wave of staged groups, in place otherwise.
	 * LZ4E_B200_DVARIANT / _DGROUP / _DSTAGE (or lz4e_b200_ctx_tune) override.
	 */
	const uint32_t resident8 = (uint32_t)ctx->sm_count * 7u * LZ4E_D_WARPS * 4u;
	const uint32_t lanes_all = (uint32_t)ctx->sm_count * 24u * 32u;
	const uint32_t stage_wave = (uint32_t)ctx->sm_count * 3u * LZ4E_D_WARPS * 4u;
	const bool can_stage = max_out_bytes <= 4096u + 64u;
	const uint32_t wave_min = ctx->wave_min >= 0 ? (uint32_t)ctx->wave_min : 0u;

	if (small && nblocks >= lanes_all) {
		group = 1;
		variant = LZ4E_DV_OLD;
		stage = 0;
	} else if (small) {
		group = 8;
		stage = can_stage && (!dst_linear || nblocks <= stage_wave);
		variant = stage ? LZ4E_DV_GROUP_PIPE : LZ4E_DV_GROUP;
	} else if (nblocks >= wave_min) {
		/* one warp per block, 32 sequences per parse/copy wave
		 * (64 KiB x 16 segments, 32768 blocks: 227 vs 168 GB/s) */
		group = 32;
		variant = LZ4E_DV_WAVE;
		stage = 0;
	} else {
		group = nblocks < resident8 ? 16 : 8;
		variant = LZ4E_DV_GROUP_PIPE;
		stage = 0;
	}
	if (ctx->dvariant >= 0)
		variant = ctx->dvariant;
	if (ctx->dgroup)
		group = ctx->dgroup;
	if (ctx->dstage == 0)
		stage = 0;
	if (ctx->dstage == 1 && can_stage)
		stage = 1;
	if (variant == LZ4E_DV_WAVE) {
		group = 32; /* one warp per block */
		stage = 0;
	} else if (group == 1)
		stage = 0;
	else if (variant != LZ4E_DV_OLD && group != 4 && group != 8 && group != 16)
		group = 8;
	cudaSetDevice(ctx->device);
	std::lock_guard<std::mutex> launch_lock(g_launch_mu);
	s = pick(ctx, stream);
	a.src_bvecs = src_bvecs;
	a.src_iters = src_iters;
	a.dst_bvecs = dst_bvecs;
	a.dst_iters = dst_iters;
	a.rets = rets;
	a.faults = ctx->d_ctl + CTL_FAULTS;
	a.nblocks = nblocks;
	a.stage_cap = stage ? (((max_out_bytes + 15u) & ~15u) + 16u) : 0u;
	a.scratch = nullptr;
	a.scratch_stride = 0;
	a.in_scratch = nullptr;
	a.in_stride = 0;
	a.order = nullptr;
	a.wave_serial = ctx->wave_serial > 0 ? (uint32_t)ctx->wave_serial : 0u;
	if (variant != LZ4E_DV_OLD || group == 1) {
		/* hand the blocks out by decreasing estimated cost (lz4e_order.cu) */
		if (ctx->dorder > 0 && nblocks >= (uint32_t)ctx->dorder) {
			uint32_t *ord = (uint32_t *)stream_scratch(
				ctx, s, SCRATCH_ORDER,
				((size_t)nblocks + 2 * LZ4E_ORDER_BUCKETS) * sizeof(uint32_t));

			if (!ord) {
				set_err(ctx, "decompress work queue: out of device memory");
				return LZ4E_B200_ENOMEM;
			}
			if (lz4e_launch_order(src_iters, dst_iters, nblocks, ord + nblocks, ord, s)) {
				set_err(ctx, "order launch failed: %s",
					cudaGetErrorString(cudaGetLastError()));
				return LZ4E_B200_EIO;
			}
			ctx->launches += 2;
			a.order = ord;
		}
	}
	if (group == 1) {
		/* lane-per-block decoder: one output slot per resident lane unless
		 * every destination is contiguous, one input slot unless every
		 * source is */
		const uint32_t ostride = dst_linear ? 0u : ((max_out_bytes + 255u) & ~255u);
		const uint32_t istride =
			(flags & LZ4E_B200_DEV_SRC_LINEAR) ? 0u : ((max_in_bytes + 255u) & ~255u);
		const size_t per_cta = (size_t)LZ4E_D_WARPS * 32 * ((size_t)ostride + istride);

		stage = 0;
		a.stage_cap = 0;
		if (per_cta) {
			int grid = lz4e_decompress_max_grid(1, 0, 0, ctx->sm_count, LZ4E_DV_OLD);
			const size_t budget = (size_t)6 << 30;
			const long long want =
				((long long)nblocks + LZ4E_D_WARPS * 32 - 1) / (LZ4E_D_WARPS * 32);

			if (grid < 1) {
				set_err(ctx, "decompress occupancy query failed");
				return LZ4E_B200_EIO;
			}
			if (grid > want)
				grid = (int)want;
			if ((size_t)grid * per_cta > budget) {
				grid = (int)(budget / per_cta);
				if (grid < 1)
					grid = 1;
			}
			grid_limit = grid;
			uint8_t *scr = stream_scratch(ctx, s, SCRATCH_DECOMPRESS, (size_t)grid * per_cta);

			if (!scr) {
				set_err(ctx, "decompress scratch: out of device memory");
				return LZ4E_B200_ENOMEM;
			}
			if (ostride) {
				a.scratch = scr;
				a.scratch_stride = ostride;
			}
			if (istride) {
				a.in_scratch = scr + (size_t)grid * LZ4E_D_WARPS * 32 * ostride;
				a.in_stride = istride;
			}
		}
	} else if (!stage && !dst_linear &&
		   !(variant != LZ4E_DV_OLD && (flags & LZ4E_B200_DEV_DST_SEGS32))) {
		/* one linear slot per resident group: the round-1 decoder decodes
		 * every scattered block there; the group decoder only blocks whose
		 * destination has more segments than its map holds */
		const uint32_t stride = (max_out_bytes + 255u) & ~255u;
		int grid = lz4e_decompress_max_grid(group, 0, 0, ctx->sm_count, variant);
		const size_t per_cta = (size_t)LZ4E_D_WARPS * (32 / group) * stride;
		const size_t budget = (size_t)2 << 30;
		const long long want = ((long long)nblocks + LZ4E_D_WARPS * (32 / group) - 1) /
				       (LZ4E_D_WARPS * (32 / group));

		if (grid < 1) {
			set_err(ctx, "decompress occupancy query failed");
			return LZ4E_B200_EIO;
		}
		if (grid > want)
			grid = (int)want;
		if ((size_t)grid * per_cta > budget) {
			grid = (int)(budget / per_cta);
			if (grid < 1)
				grid = 1;
		}
		grid_limit = grid;
		a.scratch = stream_scratch(ctx, s, SCRATCH_DECOMPRESS, (size_t)grid * per_cta);
		if (!a.scratch) {
			set_err(ctx, "decompress scratch: out of device memory");
			return LZ4E_B200_ENOMEM;
		}
		a.scratch_stride = stride;
	}
	if (group == 1 && ctx->lane_per_sm > 0) {
		/* fewer resident lanes: each keeps more of its block in L2 */
		const int cap = ctx->lane_per_sm * ctx->sm_count;

		if (grid_limit == 0 || grid_limit > cap)
			grid_limit = cap;
	}
	a.work_counter = next_counter(ctx, s);
	if (lz4e_launch_decompress(&a, group, stage, variant, ctx->sm_count, s, grid_limit,
				   nullptr)) {
		set_err(ctx, "decompress launch failed: %s",
			cudaGetErrorString(cudaGetLastError()));
		return LZ4E_B200_EIO;
	}
	ctx->launches++;
	return LZ4E_B200_OK;
}

int lz4e_b200_copy2d_h2d(lz4e_b200_ctx *ctx, void *stream, void *dst_dev, size_t dpitch,
			 const void *src_host, size_t spitch, size_t width, size_t height)
{
	cudaSetDevice(ctx->device);
	CU(ctx, cudaMemcpy2DAsync(dst_dev, dpitch, src_host, spitch, width, height,
				  cudaMemcpyHostToDevice, pick(ctx, stream)));
	ctx->h2d_bytes += width * height;
	return LZ4E_B200_OK;
}

int lz4e_b200_xfer_dev(lz4e_b200_ctx *ctx, void *stream,
		       const struct lz4e_b200_dev_bvec *bvecs,
		       const struct lz4e_b200_dev_iter *iters, const int32_t *lens,
		       uint32_t nblocks, int to_host)
{
	return lz4e_b200_xfer_skip_dev(ctx, stream, bvecs, iters, lens, nblocks, to_host, 0);
}

int lz4e_b200_xfer_skip_dev(lz4e_b200_ctx *ctx, void *stream,
			    const struct lz4e_b200_dev_bvec *bvecs,
			    const struct lz4e_b200_dev_iter *iters, const int32_t *lens,
			    uint32_t nblocks, int to_host, uint32_t skip)
{
	lz4e_xfer_args a;

	if (!ctx)
		return LZ4E_B200_ENODEV;
	if (nblocks == 0)
		return LZ4E_B200_OK;
	cudaSetDevice(ctx->device);
	a.bvecs = bvecs;
	a.iters = iters;
	a.lens = lens;
	a.moved = ctx->d_moved + (to_host ? 1 : 0);
	a.nblocks = nblocks;
	a.skip = skip;
	a.to_host = to_host ? 1 : 0;
	a.narenas = 0;
	{
		std::lock_guard<std::mutex> lk(ctx->mu);

		for (lz4e_b200_arena *ar : ctx->arenas) {
			if (a.narenas == LZ4E_XFER_MAX_ARENAS)
				break;
			a.ar[a.narenas].lo = (uint64_t)(uintptr_t)ar->dev;
			a.ar[a.narenas].hi = (uint64_t)(uintptr_t)ar->dev + ar->npages * PAGE_SIZE;
			a.ar[a.narenas].delta =
				(uint64_t)(uintptr_t)ar->host - (uint64_t)(uintptr_t)ar->dev;
			a.narenas++;
		}
	}
	for (int i = a.narenas; i < LZ4E_XFER_MAX_ARENAS; i++)
		a.ar[i].lo = a.ar[i].hi = a.ar[i].delta = 0;
	if (a.narenas == 0) {
		a.narenas = 1; /* ar[0] is an empty range: nothing matches */
	}
	if (lz4e_launch_xfer(&a, ctx->sm_count, pick(ctx, stream))) {
		set_err(ctx, "transfer kernel launch failed: %s",
			cudaGetErrorString(cudaGetLastError()));
		return LZ4E_B200_EIO;
	}
	ctx->launches++;
	return LZ4E_B200_OK;
}

int lz4e_b200_traffic(lz4e_b200_ctx *ctx, uint64_t *h2d_bytes, uint64_t *d2h_bytes)
{
	unsigned long long m[2] = { 0, 0 };

	if (!ctx)
		return LZ4E_B200_ENODEV;
	cudaSetDevice(ctx->device);
	CU(ctx, cudaMemcpy(m, ctx->d_moved, sizeof(m), cudaMemcpyDeviceToHost));
	if (h2d_bytes)
		*h2d_bytes = ctx->h2d_bytes.load() + m[0];
	if (d2h_bytes)
		*d2h_bytes = ctx->d2h_bytes.load() + m[1];
	return LZ4E_B200_OK;
}

void *lz4e_ctx_attach(lz4e_b200_ctx *ctx, int slot, size_t bytes,
		      void (*dtor)(lz4e_b200_ctx *, void *))
{
	std::lock_guard<std::mutex> lk(ctx->mu);

	for (auto &at : ctx->attached)
		if (at.slot == slot)
			return at.p;
	if (!bytes)
		return nullptr;
	void *p = calloc(1, bytes);

	if (!p)
		return nullptr;
	ctx->attached.push_back({ slot, p, dtor });
	return p;
}

void lz4e_b200_shard_range(uint64_t nblocks, uint32_t rank, uint32_t world,
			   uint64_t *first, uint64_t *count)
{
	if (world == 0)
		world = 1;
	const uint64_t lo = nblocks * rank / world;
	const uint64_t hi = nblocks * (rank + 1ull) / world;

	*first = lo;
	*count = hi - lo;
}

} /* extern "C" */

/* ---- compress without allocations ---------------------------------------- */
/*
 * Scratch of lz4e_b200_compress_dev_static (include/lz4e_b200.h): one array in
 * the library's device image, shared by every call on a device.
 *   [0, 4)      work-queue counter of the persistent warps (cleared per call)
 *   [256, end)  input slots of the global-input compress kernel, one per
 *               resident warp; the last 256 bytes stay free (the encoder's
 *               unguarded word loads may read a little past a block)
 * In host code `g_static_scratch` is only a handle for the symbol: the device
 * address comes from cudaGetSymbolAddress, once per device, with that device
 * current.
 */
#define STATIC_HEAD 256u
#define STATIC_MAX_DEVICES 64
static __device__ __align__(256) uint8_t g_static_scratch[LZ4E_B200_STATIC_SCRATCH_BYTES];
static uint8_t *g_static_ptr[STATIC_MAX_DEVICES];
static std::mutex g_static_mu;

/* device address of the array on ctx's device (which is current) */
static uint8_t *static_scratch(lz4e_b200_ctx *ctx)
{
	std::lock_guard<std::mutex> lk(g_static_mu);

	if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) {
		set_err(ctx, "static scratch: device %d outside 0..%d", ctx->device,
			STATIC_MAX_DEVICES - 1);
		return nullptr;
	}
	if (!g_static_ptr[ctx->device]) {
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);

		if (e != cudaSuccess || !p) {
			set_err(ctx, "cudaGetSymbolAddress(static scratch) failed: %s",
				cudaGetErrorString(e));
			return nullptr;
		}
		g_static_ptr[ctx->device] = (uint8_t *)p;
	}
	return g_static_ptr[ctx->device];
}

extern "C" int lz4e_b200_compress_dev_static(lz4e_b200_ctx *ctx, void *stream,
					     const struct lz4e_b200_dev_bvec *src_bvecs,
					     const struct lz4e_b200_dev_iter *src_iters,
					     const struct lz4e_b200_dev_bvec *dst_bvecs,
					     const struct lz4e_b200_dev_iter *dst_iters,
					     uint32_t nblocks, uint32_t max_in_bytes,
					     uint32_t flags, int32_t *rets,
					     struct lz4e_b200_dev_end *ends)
{
	lz4e_compress_args a;
	cudaStream_t s;
	uint8_t *scratch;
	int grid = 0;
	size_t smem = 0;

	if (!ctx)
		return LZ4E_B200_ENODEV;
	if (nblocks == 0)
		return LZ4E_B200_OK;
	if (max_in_bytes > LZ4E_B200_STATIC_MAX_BLOCK) {
		set_err(ctx, "lz4e_b200_compress_dev_static: max_in_bytes %u is above the %u "
			     "bytes its fixed %u-byte scratch array serves; use "
			     "lz4e_b200_compress_dev",
			max_in_bytes, (unsigned)LZ4E_B200_STATIC_MAX_BLOCK,
			(unsigned)LZ4E_B200_STATIC_SCRATCH_BYTES);
		return LZ4E_B200_EINVAL;
	}
	cudaSetDevice(ctx->device);
	scratch = static_scratch(ctx);
	if (!scratch)
		return LZ4E_B200_EIO;
	std::lock_guard<std::mutex> launch_lock(g_launch_mu);
	s = pick(ctx, stream);
	a.src_bvecs = src_bvecs;
	a.src_iters = src_iters;
	a.dst_bvecs = dst_bvecs;
	a.dst_iters = dst_iters;
	a.rets = rets;
	a.ends = ends;
	a.work_counter = (uint32_t *)scratch;
	a.faults = ctx->d_ctl + CTL_FAULTS; /* the context's own word, made at create time */
	a.nblocks = nblocks;
	a.order = nullptr; /* the size-ordered queue needs a buffer per batch */
	a.tma = 0;
	a.table_bytes = (flags & LZ4E_B200_DEV_NO_BYU16) ? 8192u : 16384u;
	a.in_cap = (max_in_bytes + 15u) & ~15u;
	if (a.in_cap < 16)
		a.in_cap = 16;
	{
		const uint32_t bound = max_in_bytes + max_in_bytes / 255u + 16u;

		a.out_cap = bound <= 4352u ? ((bound + 15u) & ~15u) + 16u : 1024u;
		if (a.out_cap < 1024u)
			a.out_cap = 1024u;
		if (flags & LZ4E_B200_DEV_DST_LINEAR)
			a.out_cap = 64u;
	}
	const int direct = (flags & LZ4E_B200_DEV_DST_LINEAR) ? 1 : 0;

	a.in_scratch = nullptr;
	a.in_stride = 0;
	if (lz4e_compress_plan(&a, direct, 1, ctx->sm_count, ctx->gin_per_sm, &grid, &smem)) {
		set_err(ctx, "compress launch plan failed: %s",
			cudaGetErrorString(cudaGetLastError()));
		return LZ4E_B200_EIO;
	}
	a.in_stride = ((a.in_cap + 63u) & ~255u) + 256u;
	{
		/* no more resident warps than slots in the array */
		const size_t slots =
			((size_t)LZ4E_B200_STATIC_SCRATCH_BYTES - STATIC_HEAD - 256u) / a.in_stride;

		if ((size_t)grid * a.warps > slots)
			grid = (int)(slots / a.warps);
		if (grid < 1) {
			set_err(ctx, "static scratch holds %zu slots, one CTA needs %u", slots,
				a.warps);
			return LZ4E_B200_EINVAL;
		}
	}
	a.in_scratch = scratch + STATIC_HEAD;
	/* the array keeps what the last call left: the counter must start at 0 */
	CU(ctx, cudaMemsetAsync(scratch, 0, STATIC_HEAD, s));
	if (lz4e_launch_compress(&a, direct, 1, grid, smem, s)) {
		set_err(ctx, "compress launch failed: %s", cudaGetErrorString(cudaGetLastError()));
		return LZ4E_B200_EIO;
	}
	ctx->launches++;
	return LZ4E_B200_OK;
}
