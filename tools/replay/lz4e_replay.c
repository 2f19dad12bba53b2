/*
 * tools/replay/lz4e_replay.c — BASELINE.json configs[4] / SURVEY.md row f2:
 * userspace replay of the lz4e_bdev request path through the GPU library.
 *
 * A queue of `qd` outstanding bios (fio-like: random 4 KiB-aligned offsets,
 * sizes uniform in [min,max] KiB in 4 KiB steps, `read_pct` % reads) is
 * submitted batch by batch to the userspace proxy device of
 * include/lz4e_chunk.h: every WRITE bio is compressed as one block from its own
 * bio_vecs (LZ4E_compress_default semantics), decompressed again and the
 * decompressed bytes stored (lz4e_bdev/lz4e_req.c:144-213 of the reference);
 * READs pass through (:53-75).  Reports GB/s and p50/p99 batch latency, checks
 * every read against a shadow copy of the device and the statistics counters
 * (lz4e_bdev/lz4e_stats.c:39-52).  With --oracle <liblz4e_ref.so> every write
 * bio is also compressed by the reference's own C code and the total compressed
 * size compared with what the GPU path produced (test use only).
 *
 * Build: make -C tools/replay      Run: tools/replay/lz4e_replay --help
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <errno.h>
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
#include <stdio.h>
#include <stdlib.h>
#include <pthread.h>
#include <string.h>
#include <time.h>

#include "lz4e.h"
#include "lz4e_b200.h"
#include "lz4e_chunk.h"
#include "../synth/lz4e_synth.h" /* benchmark data generator (tools/synth) */

struct opts {
	unsigned qd, batches, min_kib, max_kib, read_pct, seed, capacity_mib, device;
	int arena_bios, json;
	unsigned devices; /* > 1: one submitter thread per GPU 0 .. devices-1 in this process */
	const char *oracle;
	unsigned warmup; /* batches run and checked before the timed ones, not counted */
};

static __thread uint32_t rng_state; /* one request stream per submitter thread */
static uint32_t rnd(void)
{
	rng_state = rng_state * 1664525u + 1013904223u;
	return rng_state >> 8;
}

static double now_ms(void)
{
	struct timespec ts;

	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static int cmp_double(const void *a, const void *b)
{
	const double x = *(const double *)a, y = *(const double *)b;

	return x < y ? -1 : x > y;
}

static double pct(double *v, size_t n, double p)
{
	size_t i = (size_t)(p * (double)(n - 1) + 0.5);

	return n ? v[i < n ? i : n - 1] : 0.0;
}

/* one queue slot: a bio over up to max_pages pages of its own */
struct slot {
	struct bio bio;
	struct bio_vec *vecs;
	struct page *pages;
	unsigned int is_read;
};

static void bio_fill_pages(struct slot *s, unsigned int bytes, uint64_t pos)
{
	unsigned int left = bytes, i = 0;

	while (left) {
		const unsigned int take = left < PAGE_SIZE ? left : (unsigned int)PAGE_SIZE;

		s->vecs[i].bv_page = s->pages + i;
		s->vecs[i].bv_len = take;
		s->vecs[i].bv_offset = 0;
		left -= take;
		i++;
	}
	s->bio.bi_io_vec = s->vecs;
	s->bio.bi_vcnt = (unsigned short)i;
	s->bio.bi_iter.bi_sector = pos / 512u;
	s->bio.bi_iter.bi_size = bytes;
	s->bio.bi_iter.bi_idx = 0;
	s->bio.bi_iter.bi_bvec_done = 0;
}

/* a crash inside the library should say where */
static void on_fatal(int sig)
{
	void *bt[48];
	const int n = backtrace(bt, 48);

	fprintf(stderr, "lz4e_replay: signal %d\n", sig);
	backtrace_symbols_fd(bt, n, 2);
	_exit(128 + sig);
}

typedef int (*ref_compress_fn)(const struct bio_vec *, struct bio_vec *, struct bvec_iter *,
			       struct bvec_iter *, void *);

static void usage(void)
{
	puts("lz4e_replay [--qd 256] [--batches 64] [--warmup 0] [--min-kib 4] [--max-kib 128]\n"
	     "            [--read-pct 30] [--capacity-mib 1024] [--seed 1] [--device 0]\n"
	     "            [--devices N] [--arena-bios] [--oracle liblz4e_ref.so] [--json]");
}

static int parse(int argc, char **argv, struct opts *o)
{
	int i;

	for (i = 1; i < argc; i++) {
		const char *a = argv[i];
		const char *v = i + 1 < argc ? argv[i + 1] : NULL;

		if (!strcmp(a, "--arena-bios")) {
			o->arena_bios = 1;
		} else if (!strcmp(a, "--json")) {
			o->json = 1;
		} else if (!strcmp(a, "--help")) {
			usage();
			exit(0);
		} else if (!v) {
			return -1;
		} else if (!strcmp(a, "--qd")) {
			o->qd = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--batches")) {
			o->batches = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--warmup")) {
			o->warmup = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--min-kib")) {
			o->min_kib = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--max-kib")) {
			o->max_kib = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--read-pct")) {
			o->read_pct = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--capacity-mib")) {
			o->capacity_mib = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--seed")) {
			o->seed = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--device")) {
			o->device = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--devices")) {
			o->devices = (unsigned)atoi(v), i++;
		} else if (!strcmp(a, "--oracle")) {
			o->oracle = v, i++;
		} else {
			return -1;
		}
	}
	if (o->min_kib < 4 || o->max_kib < o->min_kib || (o->min_kib & 3) || (o->max_kib & 3) ||
	    o->max_kib > 1024 || !o->qd || !o->batches)
		return -1;
	return 0;
}

struct result {
	int rc;
	double gbs, wgbs, p50, p99, pmax;
	uint64_t wreqs, rreqs, wbytes, rbytes, cbytes, obytes, launches;
	unsigned failed, bad;
	int digest; /* -1 no oracle, 0 differs, 1 equal */
};

/* one submitter: its own context, chunk pool, stand-in device and request stream */
static int run(struct opts o, struct result *res)
{
	lz4e_b200_ctx *ctx = NULL;
	lz4e_b200_arena *arena = NULL;
	struct lz4e_dev *dev;
	struct slot *slots;
	struct bio **wb, **rb;
	int *wst, *rst;
	uint8_t *shadow, *pages_mem = NULL;
	double *lat, t_all = 0.0;
	uint64_t capacity, wbytes = 0, rbytes = 0, wreqs = 0, rreqs = 0, wvecs = 0, rvecs = 0;
	uint64_t timed_w = 0, timed_r = 0; /* bytes of the timed batches */
	uint64_t oracle_bytes = 0, oracle_digest = 0, launches0;
	unsigned int max_pages, b, i, bad = 0, failed = 0;
	ref_compress_fn ref_compress = NULL;
	void *ref_wrk = NULL;
	struct lz4e_stats ws, rs;

	memset(res, 0, sizeof(*res));
	res->rc = 3;
	rng_state = o.seed * 2654435761u + 12345u;
	capacity = (uint64_t)o.capacity_mib << 20;
	max_pages = o.max_kib / 4;
	if (lz4e_b200_ctx_create((int)o.device, &ctx) || !ctx) {
		fprintf(stderr, "lz4e_replay: no CUDA device %u: %s\n", o.device,
			lz4e_b200_strerror(NULL));
		return 3; /* no CPU fallback */
	}
	/* this thread's chunk and request-path calls run on this context */
	lz4e_chunk_bind(ctx);
	/* chunk pool: src + dst buffer per outstanding write, page-rounded */
	if (lz4e_chunk_setup(ctx, (size_t)o.qd * (2 * max_pages + 3) + 64)) {
		fprintf(stderr, "lz4e_replay: chunk pool: %s\n", lz4e_b200_strerror(ctx));
		return 3;
	}
	if (o.oracle) {
		void *h = dlopen(o.oracle, RTLD_NOW | RTLD_LOCAL);
		int (*szp)(void) = h ? (int (*)(void))dlsym(h, "oracle_sizeof_page") : NULL;
		int (*szv)(void) = h ? (int (*)(void))dlsym(h, "oracle_sizeof_bio_vec") : NULL;

		ref_compress = h ? (ref_compress_fn)dlsym(h, "LZ4E_compress_default") : NULL;
		if (!ref_compress || !szp || !szv || szp() != (int)sizeof(struct page) ||
		    szv() != (int)sizeof(struct bio_vec)) {
			fprintf(stderr, "lz4e_replay: cannot use oracle %s\n", o.oracle);
			return 2;
		}
		ref_wrk = malloc(LZ4E_MEM_COMPRESS);
	}
	dev = lz4e_dev_create(capacity);
	shadow = calloc(1, capacity);
	slots = calloc(o.qd, sizeof(*slots));
	wb = calloc(o.qd, sizeof(*wb));
	rb = calloc(o.qd, sizeof(*rb));
	wst = calloc(o.qd, sizeof(*wst));
	rst = calloc(o.qd, sizeof(*rst));
	lat = calloc(o.batches, sizeof(*lat));
	if (!dev || !shadow || !slots || !wb || !rb || !wst || !rst || !lat)
		return 3;
	/* bio pages: pageable memory (what a submitter's page cache looks like to
	 * the library: bounced) or pages of a registered arena (bulk copies) */
	if (o.arena_bios) {
		if (lz4e_b200_arena_create(ctx, (size_t)o.qd * max_pages, &arena)) {
			fprintf(stderr, "lz4e_replay: arena: %s\n", lz4e_b200_strerror(ctx));
			return 3;
		}
		pages_mem = (uint8_t *)lz4e_b200_arena_pages(arena);
	} else if (posix_memalign((void **)&pages_mem, PAGE_SIZE,
				  (size_t)o.qd * max_pages * PAGE_SIZE)) {
		return 3;
	}
	for (i = 0; i < o.qd; i++) {
		slots[i].pages = (struct page *)pages_mem + (size_t)i * max_pages;
		slots[i].vecs = calloc(max_pages, sizeof(struct bio_vec));
		slots[i].bio.bi_max_vecs = (unsigned short)max_pages;
	}
	launches0 = lz4e_b200_kernel_launches(ctx);

	for (b = 0; b < o.warmup + o.batches; b++) {
		size_t nw = 0, nr = 0, k;
		const int timed = b >= o.warmup;
		double t0;

		/* ---- fill the queue ---- */
		for (i = 0; i < o.qd; i++) {
			struct slot *s = &slots[i];
			const unsigned int steps = (o.max_kib - o.min_kib) / 4 + 1;
			const unsigned int bytes = (o.min_kib + 4 * (rnd() % steps)) * 1024u;
			const uint64_t pos = (uint64_t)(rnd() % ((capacity - bytes) / PAGE_SIZE + 1)) *
					     PAGE_SIZE;

			/* the first batch only writes, so reads find data */
			s->is_read = b > 0 && rnd() % 100 < o.read_pct;
			bio_fill_pages(s, bytes, pos);
			s->bio.bi_opf = s->is_read ? 0 : 1; /* REQ_OP_READ / REQ_OP_WRITE */
			if (s->is_read) {
				memset(s->pages, 0xA5, bytes);
				rb[nr++] = &s->bio;
				rbytes += bytes, rreqs++, rvecs += s->bio.bi_vcnt;
				if (timed)
					timed_r += bytes;
			} else {
				unsigned int off;

				/* content: one synthetic class per 4 KiB..64 KiB piece */
				for (off = 0; off < bytes;) {
					unsigned int piece = 4096u << (rnd() % 5);

					if (piece > bytes - off)
						piece = bytes - off;
					lz4e_synth_block((uint8_t *)s->pages + off, piece,
								  rnd() & 3, rnd());
					off += piece;
				}
				wb[nw++] = &s->bio;
				wbytes += bytes, wreqs++, wvecs += s->bio.bi_vcnt;
				if (timed)
					timed_w += bytes;
				if (ref_compress) {
					/* the reference on the same bio, into scratch pages */
					const unsigned int cap = LZ4E_COMPRESSBOUND(bytes);
					const unsigned int np = (cap + PAGE_SIZE - 1) / PAGE_SIZE;
					struct page *dp = aligned_alloc(PAGE_SIZE, (size_t)np * PAGE_SIZE);
					struct bio_vec dv = { dp, np * (unsigned int)PAGE_SIZE, 0 };
					struct bvec_iter si = s->bio.bi_iter, di = { 0, cap, 0, 0 };
					const int r = ref_compress(s->vecs, &dv, &si, &di, ref_wrk);

					if (r <= 0)
						bad++;
					else
						oracle_digest += lz4e_fnv1a64(dp, (size_t)r);
					oracle_bytes += (uint64_t)r;
					free(dp);
				}
			}
		}
		/* ---- submit: one GPU batch for all writes, then the reads ---- */
		t0 = now_ms();
		lz4e_dev_submit_writes(dev, wb, nw, wst);
		lz4e_dev_submit_reads(dev, rb, nr, rst);
		if (timed) {
			lat[b - o.warmup] = now_ms() - t0;
			t_all += lat[b - o.warmup];
		}
		/* ---- check ---- */
		for (k = 0; k < nw; k++) {
			const struct bio *w = wb[k];

			if (wst[k]) {
				failed++;
				continue;
			}
			memcpy(shadow + w->bi_iter.bi_sector * 512u,
			       page_address(w->bi_io_vec[0].bv_page), w->bi_iter.bi_size);
		}
		for (k = 0; k < nr; k++) {
			const struct bio *r = rb[k];

			if (rst[k] || memcmp(shadow + r->bi_iter.bi_sector * 512u,
					     page_address(r->bi_io_vec[0].bv_page),
					     r->bi_iter.bi_size))
				bad++;
		}
	}

	/* ---- statistics counters, lz4e_stats.c:39-52 ---- */
	lz4e_dev_stats(dev, 1, &ws);
	lz4e_dev_stats(dev, 0, &rs);
	if (ws.reqs_total != wreqs || ws.reqs_failed != failed || ws.data_in_bytes + 0 != wbytes ||
	    ws.vec_count != wvecs || rs.reqs_total != rreqs || rs.data_in_bytes != rbytes ||
	    rs.vec_count != rvecs || rs.reqs_failed != 0) {
		fprintf(stderr, "lz4e_replay: statistics differ: writes %llu/%llu reqs, %llu/%llu bytes, "
				"%llu/%llu vecs; reads %llu/%llu reqs\n",
			(unsigned long long)ws.reqs_total, (unsigned long long)wreqs,
			(unsigned long long)ws.data_in_bytes, (unsigned long long)wbytes,
			(unsigned long long)ws.vec_count, (unsigned long long)wvecs,
			(unsigned long long)rs.reqs_total, (unsigned long long)rreqs);
		bad++;
	}
	if (ref_compress && !failed && oracle_digest != lz4e_dev_compressed_digest(dev)) {
		/* per-bio BYTES: the sum of per-bio FNV-1a-64 digests must agree */
		fprintf(stderr, "lz4e_replay: compressed bytes differ from the reference's "
				"(digest %016llx vs %016llx)\n",
			(unsigned long long)lz4e_dev_compressed_digest(dev),
			(unsigned long long)oracle_digest);
		bad++;
	}
	if (ref_compress && oracle_bytes != lz4e_dev_compressed_bytes(dev)) {
		fprintf(stderr, "lz4e_replay: compressed %llu bytes, the reference %llu\n",
			(unsigned long long)lz4e_dev_compressed_bytes(dev),
			(unsigned long long)oracle_bytes);
		bad++;
	}
	qsort(lat, o.batches, sizeof(*lat), cmp_double);
	res->gbs = (double)(timed_w + timed_r) / (t_all * 1e-3) / 1e9;
	res->wgbs = (double)timed_w / (t_all * 1e-3) / 1e9;
	res->p50 = pct(lat, o.batches, 0.50);
	res->p99 = pct(lat, o.batches, 0.99);
	res->pmax = lat[o.batches - 1];
	res->wreqs = wreqs, res->rreqs = rreqs, res->wbytes = wbytes, res->rbytes = rbytes;
	res->cbytes = lz4e_dev_compressed_bytes(dev);
	res->obytes = oracle_bytes;
	res->digest = !ref_compress ? -1 : oracle_digest == lz4e_dev_compressed_digest(dev);
	res->failed = failed, res->bad = bad;
	res->launches = lz4e_b200_kernel_launches(ctx) - launches0;
	for (i = 0; i < o.qd; i++)
		free(slots[i].vecs);
	lz4e_dev_destroy(dev);
	lz4e_chunk_bind(NULL);
	/* the context takes its chunk pool and arenas with it */
	if (!arena)
		free(pages_mem);
	lz4e_b200_ctx_destroy(ctx);
	free(shadow);
	res->rc = bad || failed ? 1 : 0;
	return res->rc;
}

struct job {
	struct opts o;
	struct result r;
};

static void *job_main(void *p)
{
	struct job *j = p;

	run(j->o, &j->r);
	return NULL;
}

int main(int argc, char **argv)
{
	struct opts o = { 256, 64, 4, 128, 30, 1, 1024, 0, 0, 0, 1, NULL };
	struct job jobs[16];
	pthread_t th[16];
	struct result t;
	unsigned d, n;
	double t0, wall;

	signal(SIGBUS, on_fatal);
	signal(SIGSEGV, on_fatal);
	if (parse(argc, argv, &o) || o.devices < 1 || o.devices > 16) {
		usage();
		return 2;
	}
	n = o.devices;
	for (d = 0; d < n; d++) {
		jobs[d].o = o;
		if (n > 1) {
			/* submitter d on GPU d; with fewer GPUs than submitters they
			 * share (each still has its own context and pool) */
			const int ndev = lz4e_b200_device_count();

			jobs[d].o.device = ndev > 0 ? d % (unsigned)ndev : d;
			jobs[d].o.seed = o.seed + d;
		}
	}
	t0 = now_ms();
	for (d = 1; d < n; d++)
		pthread_create(&th[d], NULL, job_main, &jobs[d]);
	job_main(&jobs[0]);
	for (d = 1; d < n; d++)
		pthread_join(th[d], NULL);
	wall = now_ms() - t0;
	/* throughput and counters add up, latencies are the slowest submitter's */
	t = jobs[0].r;
	for (d = 1; d < n; d++) {
		const struct result *r = &jobs[d].r;

		if (r->rc > t.rc)
			t.rc = r->rc;
		t.gbs += r->gbs, t.wgbs += r->wgbs;
		if (r->p50 > t.p50)
			t.p50 = r->p50;
		if (r->p99 > t.p99)
			t.p99 = r->p99;
		if (r->pmax > t.pmax)
			t.pmax = r->pmax;
		t.wreqs += r->wreqs, t.rreqs += r->rreqs, t.wbytes += r->wbytes, t.rbytes += r->rbytes;
		t.cbytes += r->cbytes, t.obytes += r->obytes, t.launches += r->launches;
		t.failed += r->failed, t.bad += r->bad;
		if (r->digest < t.digest)
			t.digest = r->digest;
	}
	if (t.rc == 3) {
		fprintf(stderr, "lz4e_replay: set-up failed (no CUDA device or out of memory)\n");
		return 3;
	}
	printf("{\"workload\": \"chunk-layer replay: QD%u, %u batches, %u-%u KiB, %u%% reads, "
	       "%s bio pages%s\", \"gbs\": %.3f, \"write_gbs\": %.3f, \"batch_ms_p50\": %.3f, "
	       "\"batch_ms_p99\": %.3f, \"batch_ms_max\": %.3f, \"write_reqs\": %llu, "
	       "\"read_reqs\": %llu, \"write_bytes\": %llu, \"read_bytes\": %llu, "
	       "\"compressed_bytes\": %llu, \"oracle_compressed_bytes\": %llu, "
	       "\"per_bio_digest_equal\": %s, \"failed\": %u, \"mismatches\": %u, "
	       "\"gpu_launches\": %llu, \"submitters\": %u, \"wall_ms\": %.1f, "
	       "\"warmup_batches\": %u}\n",
	       o.qd, o.batches, o.min_kib, o.max_kib, o.read_pct,
	       o.arena_bios ? "arena (pinned, mirrored)" : "pageable (bounced)",
	       n > 1 ? ", one submitter thread per GPU in one process" : "", t.gbs, t.wgbs, t.p50,
	       t.p99, t.pmax, (unsigned long long)t.wreqs, (unsigned long long)t.rreqs,
	       (unsigned long long)t.wbytes, (unsigned long long)t.rbytes,
	       (unsigned long long)t.cbytes, (unsigned long long)t.obytes,
	       t.digest < 0 ? "null" : t.digest ? "true" : "false", t.failed, t.bad,
	       (unsigned long long)t.launches, n, wall, o.warmup);
	return t.rc;
}
