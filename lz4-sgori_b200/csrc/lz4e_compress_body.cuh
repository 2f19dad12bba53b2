/*
 * lz4e_compress_body.cuh — one warp compresses one block.
 *
 * What it must reproduce, byte for byte: LZ4E_compress_generic
 * (lz4e/lz4e_compress.c:218-534 of the reference) as instantiated by
 * LZ4E_compress_default (:536-569): greedy LZ4 "fast", acceleration 1, noDict,
 * hash-table address mode chosen from the SOURCE SEGMENTATION by
 * LZ4E_fillBvIterSize (:184-211).
 *
 * How it is done here (none of this is in the reference):
 *
 *  - The block is gathered once from its SG segments into shared memory as a
 *    linear byte string (lz4e_copy.cuh), so the per-access bio_vec walk of
 *    lz4e_defs.h:366-402 disappears from the hot loop.
 *  - The hash table lives in shared memory, same number of entries as the
 *    reference (8192 / 4096 / 2048 for byU16 / byU32 / byU64,
 *    lz4e_compress.c:48-57) but holding 16-bit LINEAR positions: the
 *    reference's (bvec index, offset) address (lz4e_defs.h:641-686) is a
 *    bijection on positions and "raw == 0" means position 0 in both.
 *  - The reference's search loop (:292-336) visits positions p_0, p_1, ... with
 *    a fixed step schedule and, at each one, reads table[h], writes table[h]
 *    and compares 4 bytes.  A visit that does not match changes nothing but
 *    the table.  So 32 consecutive visits are evaluated SPECULATIVELY by the 32
 *    lanes; __match_any_sync resolves visits of one batch that hit the same
 *    table slot (a later visit must see the earlier visit's position), the
 *    first matching lane wins, and only visits up to and including the winner
 *    are committed to the table — exactly the serial order.
 *  - The two extra table operations after every match ("fill table" at ip-2,
 *    :461-464, and "test next position" at ip, :467-493) are lanes -2 and -1 of
 *    the next batch.
 *  - Catch-up (:339-349) and LZ4E_count (lz4e_defs.h:587-636) are
 *    ballot/ffs reductions over per-lane compares.
 *  - A sequence is emitted only when literal length AND match length are
 *    known, so the output is append-only; it is staged in shared memory and
 *    flushed to the destination SG segments in coalesced pieces.
 */
#ifndef LZ4E_COMPRESS_BODY_CUH
#define LZ4E_COMPRESS_BODY_CUH

#include "../../include/lz4e_b200.h"
#include "lz4e_copy.cuh"
#include "lz4e_simt.h"

#define LZ4E_K_MAX_INPUT 0x7E000000u /* lz4e.h:24 */
#define LZ4E_K_TABLE_BYTES 16384u    /* LZ4E_MEMORY_USAGE 14, lz4e.h:11 */
#define LZ4E_K_PAD 48u		     /* slack after each staging area */
#define LZ4E_K_BAR 16u		     /* the warp's mbarrier (TMA staging of the input) */

#define LZ4E_FAULT_RANGE 1u /* a block exceeded the limits given at launch */

enum { LZ4E_TT_U16 = 1, LZ4E_TT_U32 = 3, LZ4E_TT_U64 = 7 }; /* lz4e_defs.h:689 */

struct lz4e_compress_args {
	const struct lz4e_b200_dev_bvec *src_bvecs;
	const struct lz4e_b200_dev_iter *src_iters;
	const struct lz4e_b200_dev_bvec *dst_bvecs;
	const struct lz4e_b200_dev_iter *dst_iters;
	int32_t *rets;
	struct lz4e_b200_dev_end *ends; /* may be NULL */
	uint32_t *work_counter;		/* zeroed before launch */
	const uint32_t *order;		/* optional: block indices, most expensive first */
	uint32_t *faults;
	uint32_t nblocks;
	uint32_t table_bytes; /* 16384, or 8192 when no block can select byU16 */
	uint32_t in_cap;  /* bytes of input staging (multiple of 16) */
	uint32_t out_cap; /* bytes of output staging (multiple of 16, >= 1024 and
			   * >= in_cap / 255 + 20; 64 = written in place) */
	/* GIN kernels only: the block is gathered into a linear slot in global
	 * memory (one slot per resident warp) and read back through L1, so that
	 * shared memory holds just the hash table and large blocks no longer
	 * limit the number of resident warps */
	uint8_t *in_scratch;
	uint32_t in_stride; /* bytes per slot, >= in_cap + 64 */
	/* CTA shape, set by lz4e_compress_plan(): warps per CTA and the shared
	 * memory stride from one warp's area to the next */
	uint32_t warps;
	uint32_t warp_smem;
	/* shared-memory kernels: stage congruent segments with 1-D TMA bulk copies
	 * (cp.async.bulk + mbarrier) instead of loads/stores by the lanes */
	uint32_t tma;
};

/* shared memory needed by one warp (in_cap = 0 for the GIN kernels) */
static inline uint32_t lz4e_compress_smem_bytes(uint32_t table_bytes,
						uint32_t in_cap, uint32_t out_cap)
{
	return table_bytes + (in_cap ? in_cap + LZ4E_K_PAD : 0u) + out_cap + LZ4E_K_PAD +
	       LZ4E_K_BAR;
}

/* number of bytes advanced after k search visits: visit k uses step 1 for
 * k == 0 and (63 + k) >> 6 afterwards (searchMatchNb++ >> LZ4_SKIPTRIGGER,
 * lz4e_compress.c:296-306) */
LZ4E_DEV uint32_t lz4e_search_advance(int k)
{
	if (k <= 65)
		return (uint32_t)k;
	const uint32_t a1 = 63u + (uint32_t)k; /* A + 1 with A = 62 + k */
	const uint32_t q = a1 >> 6;

	return 1u + 32u * q * (q - 1u) + (a1 - 64u * q) * q;
}

/* hash of the bytes at a position, lz4e_compress.c:59-96; w0/w1 are the two
 * aligned words around the position, sh its bit offset in w0 */
LZ4E_DEV uint32_t lz4e_hash(uint32_t seq, uint32_t byte4, int tt)
{
	if (tt == LZ4E_TT_U32) {
		const uint64_t v = (uint64_t)seq | ((uint64_t)byte4 << 32);

		return (uint32_t)(((v << 24) * 889523592379ULL) >> 52);
	}
	return (seq * 2654435761U) >> (tt == LZ4E_TT_U16 ? 19 : 21);
}

/*
 * Word access to the gathered block.  `da` is the 4-byte aligned shared-memory
 * address just below the block's byte 0 and `t` = position + (d - da), so the
 * aligned word holding a byte is da + (t & ~3) with 32-bit arithmetic only:
 * the address keeps its shared address space (LDS, no generic 64-bit math).
 * The funnel shift takes its count modulo 32, so t << 3 is the bit offset.
 */
LZ4E_DEV uint32_t lz4e_sld32u(const uint8_t *da, uint32_t t)
{
	const uint8_t *al = da + (t & ~3u);

	return lz4e_funnel_r(lz4e_ld32a(al), lz4e_ld32a(al + 4), t << 3);
}

/* the 4 bytes at t (little endian) and the byte after them */
LZ4E_DEV uint32_t lz4e_sld_seq5(const uint8_t *da, uint32_t t, uint32_t &byte4)
{
	const uint8_t *al = da + (t & ~3u);
	const uint32_t w0 = lz4e_ld32a(al);
	const uint32_t w1 = lz4e_ld32a(al + 4);

	byte4 = (w1 >> ((t & 3u) * 8u)) & 0xffu;
	return lz4e_funnel_r(w0, w1, t << 3);
}

struct lz4e_out_stage {
	uint8_t *base;	    /* staging area (shared memory) */
	uint8_t *cur0;	    /* address holding output byte `win` */
	uint32_t win;	    /* linear output position of cur0 */
	uint32_t cap;	    /* usable bytes behind cur0 */
	lz4e_cursor dst;    /* where output byte `win` goes */
};

LZ4E_DEV void lz4e_stage_rebase(lz4e_out_stage &s, uint32_t op, uint32_t out_cap)
{
	/* keep staging and destination congruent mod 16 so the flush is vector
	 * stores */
	const uint32_t shift = (uint32_t)(s.dst.v->addr + s.dst.done) & 15u;

	s.cur0 = s.base + shift;
	s.win = op;
	s.cap = out_cap;
}

LZ4E_DEV void lz4e_stage_flush(lz4e_out_stage &s, uint32_t op, uint32_t out_cap,
			       uint32_t lane)
{
	lz4e_syncwarp(LZ4E_FULL);
	lz4e_scatter<32>(s.dst, s.cur0, op - s.win, lane);
	lz4e_syncwarp(LZ4E_FULL);
	lz4e_stage_rebase(s, op, out_cap);
}

/* length bytes of a run >= 15: (run - 15) / 255 bytes of 255, then the rest */
LZ4E_DEV void lz4e_put_run(uint8_t *p, uint32_t count, uint32_t last,
			   uint32_t lane)
{
	/* almost always one or two bytes: a single predicated store */
	if (lane < count)
		p[lane] = (lane + 1u == count) ? (uint8_t)last : (uint8_t)255;
	if (count > 32u)
		for (uint32_t i = lane + 32u; i < count; i += 32)
			p[i] = (i + 1 == count) ? (uint8_t)last : (uint8_t)255;
}

/*
 * limitedOutput failure inside a sequence (lz4e_compress.c:358-363 and
 * :425-430).  The reference returns 0 with its iterators where they were and
 * with what it had stored so far; callers that look at *dstIter after a 0 see
 * that state, so it is reproduced: everything before the sequence is flushed,
 * and for the second check (`second`) the literal length bytes, the literals
 * and the offset are stored too.  The token byte of the sequence is never
 * written by the reference on these paths and is left alone here.  Cold and
 * out of line: the encoder's emit path only tests for it.
 */
template <bool DIRECT>
LZ4E_DEV_NOINLINE void lz4e_compress_overflow(uint8_t *cur0, uint8_t *base, uint32_t win,
					      const struct lz4e_b200_dev_bvec *dv, uint32_t ddone,
					      const uint8_t *d, uint32_t op, uint32_t anchor,
					      uint32_t lit, uint32_t lit_ext, uint32_t off,
					      bool second, uint32_t lane)
{
	/* everything by value: taking the address of the encoder's staging state
	 * would move it from registers to local memory on the hot path */
	lz4e_cursor c = { dv, ddone };

	if (DIRECT) {
		uint8_t *o = cur0 + op + 1u;

		if (!second)
			return;
		if (lit_ext)
			lz4e_put_run(o, lit_ext, (lit - 15u) % 255u, lane);
		o += lit_ext;
		lz4e_copy<32>(o, d + anchor, lit, lane);
		if (lane < 2u)
			o[lit + lane] = (uint8_t)(lane ? off >> 8 : off);
		return;
	}
	lz4e_syncwarp(LZ4E_FULL);
	lz4e_scatter<32>(c, cur0, op - win, lane);
	lz4e_syncwarp(LZ4E_FULL);
	if (!second)
		return;
	lz4e_cursor_skip(c, 1u); /* the token byte */
	for (uint32_t left = lit_ext; left;) {
		const uint32_t part = left < 512u ? left : 512u;

		lz4e_put_run(base, part, part == left ? (lit - 15u) % 255u : 255u, lane);
		lz4e_syncwarp(LZ4E_FULL);
		lz4e_scatter<32>(c, base, part, lane);
		lz4e_syncwarp(LZ4E_FULL);
		left -= part;
	}
	lz4e_scatter<32>(c, d + anchor, lit, lane);
	if (lane < 2u)
		base[lane] = (uint8_t)(lane ? off >> 8 : off);
	lz4e_syncwarp(LZ4E_FULL);
	lz4e_scatter<32>(c, base, 2u, lane);
}

/*
 * The hash table in shared memory.  Plain form: one TABT per entry.  Packed
 * form (P12, blocks of at most 4 KiB in byU16 mode): positions fit 12 bits, so
 * the 8192 entries take 12 KiB instead of 16 — low bytes in [0, 8192), high
 * nibbles two to a byte in [8192, 12288) — and 14 instead of 11 warps are
 * resident per SM.  Entry count, hash and "0 = empty = position 0" are the
 * reference's (lz4e_compress.c:48-57, :98-168), so the bytes produced cannot
 * change.  Two lanes that write entries h and h ^ 1 share a nibble byte: the
 * warp-wide commits go in two passes (even entries, then odd ones).
 */
#define LZ4E_K_P12_BYTES 12288u
template <typename TABT, bool P12>
struct lz4e_table {
	uint8_t *sm;

	LZ4E_MEMBER uint32_t get(uint32_t h) const
	{
		if (P12) {
			const uint32_t lo = sm[h];
			const uint32_t hi = sm[8192u + (h >> 1)];

			return lo | (((hi >> ((h & 1u) * 4u)) & 15u) << 8);
		}
		return reinterpret_cast<const TABT *>(sm)[h];
	}
	/* one lane, or lanes that are known not to share a nibble byte */
	LZ4E_MEMBER void put(uint32_t h, uint32_t v) const
	{
		if (P12) {
			uint8_t *p = sm + 8192u + (h >> 1);
			const uint32_t sh = (h & 1u) * 4u;

			sm[h] = (uint8_t)v;
			*p = (uint8_t)((*p & ~(15u << sh)) | ((v >> 8) << sh));
		} else {
			reinterpret_cast<TABT *>(sm)[h] = (TABT)v;
		}
	}
	/* every lane with `on` writes its own entry (all h distinct) */
	LZ4E_MEMBER void commit(bool on, uint32_t h, uint32_t v) const
	{
		if (P12) {
			if (on && !(h & 1u))
				put(h, v);
			lz4e_syncwarp(LZ4E_FULL);
			if (on && (h & 1u))
				put(h, v);
		} else if (on) {
			put(h, v);
		}
	}
};

/*
 * The encoder proper for one gathered block: d[0..n) in shared memory, table
 * `tab` (not yet cleared), output staging `out_area`.  TT is the hash-table
 * address mode LZ4E_fillBvIterSize selected (lz4e_compress.c:184-211).
 * Returns the reference's return value; src_end/dst_end receive the iterator
 * end state.
 */
template <int TT, bool DIRECT, bool GIN, typename TABT, bool P12 = false>
LZ4E_DEV int32_t lz4e_compress_encode(const lz4e_compress_args &a,
				      const struct lz4e_b200_dev_iter &di,
				      const uint8_t *d, uint32_t n, uint8_t *sm,
				      uint8_t *out_area, uint32_t lane,
				      uint32_t &src_end, uint32_t &dst_end)
{
	const uint32_t cap = di.bi_size;
	const lz4e_table<TABT, P12> tab = { sm };
	int32_t ret = 0;
	/* 16-bit positions: the block is at most 64 KiB, so every candidate is
	 * within LZ4E_DISTANCE_MAX (lz4e_compress.c:326-327 can never fire) */
	constexpr bool NEAR = sizeof(TABT) == 2;

	if (GIN)
		LZ4E_ASSUME_GLOBAL(d);
	else
		LZ4E_ASSUME_SHARED(d);
	LZ4E_ASSUME_SHARED(sm);
	if (!DIRECT)
		LZ4E_ASSUME_SHARED(out_area);
	/* see lz4e_sld32u(): aligned base and the block's offset from it */
	const uint32_t dm = (uint32_t)(uintptr_t)d & 3u;
	const uint8_t *da = d - dm;

	if (GIN)
		LZ4E_ASSUME_GLOBAL(da);
	else
		LZ4E_ASSUME_SHARED(da);

	{
		const bool limited = cap < n + n / 255u + 16u; /* :553 */
		const uint32_t mflimit = n - 12u;
		const uint32_t matchlimit = n - 5u;
		/* the whole output fits the staging area: one flush at the end */
		const bool whole = DIRECT || a.out_cap >= (cap < n + n / 255u + 16u ? cap : n + n / 255u + 16u);
		/* streaming output (a window smaller than the block's bound) takes the
		 * fast chain too: the window is flushed when four more bytes do not fit */
		const bool fast_ok = !limited;
		uint32_t anchor = 0, op = 0;
		lz4e_out_stage st;

		st.base = out_area;
		st.dst = lz4e_cursor_init(a.dst_bvecs, di);
		if (DIRECT) {
			/* the destination is one contiguous segment (checked by
			 * the caller): sequences are stored where they belong,
			 * no staging, no flush */
			st.cur0 = cap ? lz4e_cursor_addr(st.dst) : out_area;
			st.win = 0;
			st.cap = 0xffffffffu;
		} else if (cap) {
			lz4e_stage_rebase(st, 0, a.out_cap);
		} else {
			st.cur0 = out_area, st.win = 0, st.cap = a.out_cap;
		}

		if (n >= 13) {
			/* the table — (1 << hashLog) entries of TABT, 16-bit positions
			 * up to 64 KiB, 32-bit ones above — was zeroed by the caller
			 * while the block's bytes were still in flight */
			const uint32_t tbytes = P12 ? LZ4E_K_P12_BYTES
						    : (TT == LZ4E_TT_U16   ? 8192u
						       : TT == LZ4E_TT_U32 ? 4096u
									   : 2048u) * (uint32_t)sizeof(TABT);

			if (tbytes > a.table_bytes) {
				/* launched for byU32-only batches, block wants byU16 */
				if (lane == 0)
					lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
				return 0;
			}

			/* first byte (:280-282): table[hash(0)] = 0 is a no-op on a
			 * zeroed table of positions; search starts at 1 */
			uint32_t p0 = 1;
			int kbase = 0;
			bool have = false; /* (ip, mp) already known, see below */
			uint32_t ip = 0, mp = 0;

			src_end = 1;
			for (;;) {
				if (!have) {
				/* ---- one speculative batch of 32 visits ---- */
				const int k = kbase + (int)lane;
				uint32_t pos, seq = 0, h = 0xffffffffu, m = 0;
				bool valid;

				if (kbase + 31 <= 64) {
					/* step is still 1 everywhere in this batch; lanes
					 * -2 / -1 are the post-match table fill at ip-2 and
					 * the "test next position" visit at ip */
					pos = p0 + (uint32_t)k - (k == -2 ? 1u : 0u);
					valid = k < 0 || pos + 1u <= mflimit;
				} else {
					const uint32_t step = (63u + (uint32_t)k) >> 6;

					pos = p0 + lz4e_search_advance(k);
					valid = pos + step <= mflimit;
				}
				if (valid) {
					uint32_t b4;

					seq = lz4e_sld_seq5(da, pos + dm, b4);
					h = lz4e_hash(seq, b4, TT);
					m = tab.get(h);
				}
				const uint32_t peers = lz4e_match_any(LZ4E_FULL, h);
				const uint32_t lower = peers & ((1u << lane) - 1u);
				const int srcl = lower ? 31 - lz4e_clz(lower) : (int)lane;
				const uint32_t mj = lz4e_shfl(LZ4E_FULL, pos, srcl);
				bool ok = false;

				if (lower)
					m = mj;
				if (valid && k != -2 && (NEAR || m + 65535u >= pos))
					ok = lz4e_sld32u(da, m + dm) == seq;

				const uint32_t okmask = lz4e_ballot(LZ4E_FULL, ok);

				if (okmask == 0) {
					const uint32_t vmask = lz4e_ballot(LZ4E_FULL, valid);

					if (vmask != LZ4E_FULL) {
						/* ran into mflimit: last literals.  *srcIter
						 * stays at the last visited search position */
						const int lastk = kbase + lz4e_popc(vmask) - 1;

						src_end = p0 + (lastk >= 0 ? lz4e_search_advance(lastk)
									   : 0u);
						break;
					}
					/* commit all 32 visits, keep searching */
					tab.commit((peers >> lane) >> 1 == 0, h, pos);
					lz4e_syncwarp(LZ4E_FULL);
					kbase += 32;
					continue;
				}

				/* ---- a match: commit visits 0..f in serial order ---- */
				const int f = lz4e_ffs(okmask) - 1;

				tab.commit((int)lane <= f &&
						   (peers & ((2u << f) - 1u) & ~((2u << lane) - 1u)) == 0,
					   h, pos);

				ip = lz4e_shfl(LZ4E_FULL, pos, f);
				mp = lz4e_shfl(LZ4E_FULL, m, f);
				lz4e_syncwarp(LZ4E_FULL); /* table writes visible */
				have = true;
				}

				/*
				 * ---- fast chain: a match that starts exactly at the anchor
				 * (no literals, no catch-up) and is shorter than 36 bytes ----
				 * This is what "test next position" (:467-493) produces over
				 * and over on compressible data.  Lane j assumes the match is
				 * j bytes past the minimum, i.e. that the next anchor is
				 * q = ip + 4 + j, and evaluates the reference's next test
				 * there (table[H(q-2)] = q-2; candidate = table[H(q)]; compare
				 * 4 bytes) from three aligned words around q-2; the table is
				 * not touched between the match decision and that test, so
				 * reading it now is exact.  The lane that guessed right hands
				 * its result over.
				 */
				bool ended = false;

				if (fast_ok && have && ip == anchor) {
				/*
				 * The input words of a round are loaded one round EARLY: as
				 * soon as the length of this match (-> next anchor) and its
				 * successor's candidate are known, i.e. before this round's
				 * validity vote, sequence bytes and table update.  The block
				 * is read-only, so the early loads cannot see a stale value;
				 * the reads may reach a few bytes into the slack behind the
				 * block like the unguarded loads of the round itself.
				 */
				uint32_t w0, w1, w2, sh, mb;
				{
					const uint32_t tq = ip + 4u + lane - 2u + dm;
					const uint8_t *al = da + (tq & ~3u);

					sh = tq << 3;
					w0 = lz4e_ld32a(al), w1 = lz4e_ld32a(al + 4), w2 = lz4e_ld32a(al + 8);
					mb = d[mp + 4u + lane];
				}
				for (;;) {
					const uint32_t fip = ip + 4u, q = fip + lane;
					const uint32_t lo = lz4e_funnel_r(w0, w1, sh); /* bytes q-2.. */
					const uint32_t hi = lz4e_funnel_r(w1, w2, sh); /* bytes q+2.. */
					const uint32_t sq = lz4e_funnel_r(lo, hi, 16u); /* bytes q..   */
					const uint32_t qh = lz4e_hash(sq, (hi >> 16) & 0xffu, TT);
					const uint32_t qh2 = lz4e_hash(lo, hi & 0xffu, TT);
					const uint32_t qm = qh2 == qh ? q - 2u : tab.get(qh);
					const bool fne = lane >= matchlimit - fip || (sq & 0xffu) != mb;
					const uint32_t fnz = lz4e_ballot(LZ4E_FULL, fne);

					if (fnz == 0)
						break; /* >= 32 more bytes: general path */
					const uint32_t ml = (uint32_t)lz4e_ffs(fnz) - 1u;
					const uint32_t na = fip + ml;
					const uint32_t off = ip - mp;
					const uint32_t cand = lz4e_sld32u(da, qm + dm);

					mp = lz4e_shfl(LZ4E_FULL, qm, (int)ml);
					{ /* next round's loads, see above */
						const uint32_t tq = na + 4u + lane - 2u + dm;
						const uint8_t *al = da + (tq & ~3u);

						sh = tq << 3;
						w0 = lz4e_ld32a(al), w1 = lz4e_ld32a(al + 4),
						w2 = lz4e_ld32a(al + 8);
						mb = d[mp + 4u + lane];
					}
					const bool qok = (NEAR || qm + 65535u >= q) && cand == sq;

					if (!whole && op - st.win + 4u > st.cap)
						lz4e_stage_flush(st, op, a.out_cap, lane);
					uint8_t *o = st.cur0 + (op - st.win);
					const bool last = na > mflimit; /* :456 */

					/* token, offset (LE16) and, from 15 on, one length byte:
					 * lane l stores byte l of one packed word */
					const uint32_t word = (ml < 15u ? ml : 15u) | (off << 8) |
							      ((ml - 15u) << 24);

					if (lane < 3u + (ml >= 15u ? 1u : 0u))
						o[lane] = (uint8_t)(word >> (8u * lane));
					op += 3u + (ml >= 15u ? 1u : 0u);
					const uint32_t okm = lz4e_ballot(LZ4E_FULL, qok);

					/* the lane that guessed the length holds the hashes of
					 * :461-464 and :467-470 for the new anchor q == na */
					if (lane == ml && !last) {
						tab.put(qh2, q - 2u);
						tab.put(qh, q);
					}
					anchor = ip = na;
					p0 = na + 1u; /* if no match follows: search from ip + 1 */
					kbase = 0;
					ended = last;
					have = !last && ((okm >> ml) & 1u);
					lz4e_syncwarp(LZ4E_FULL); /* table writes visible */
					if (!have)
						break;
				}
				}
				if (ended) {
					src_end = anchor;
					break;
				}
				if (!have)
					continue;

				/*
				 * ---- catch up (:339-349), match length (lz4e_defs.h:587-636)
				 * and the NEXT "test next position" (:461-493), all at once ----
				 * One byte per lane each way.  Lane j also assumes the match
				 * is j bytes long, i.e. that the next anchor is q = ip + 4 + j,
				 * and evaluates what the reference does there: table[H(q-2)] =
				 * q-2, candidate = table[H(q)], compare 4 bytes.  The table is
				 * not touched between the match decision and that point, so
				 * reading it now is exact; once the length is known the lane
				 * that guessed right hands over its result and the serial
				 * chain from one match to the next is a single round of loads.
				 */
				uint32_t back = 0, ml;
				uint32_t q_h = 0, q_h2 = 0, q_m = 0;
				bool q_ok = false;
				{
					const uint32_t fip = ip + 4u, fm = mp + 4u;
					const uint32_t limit = matchlimit - fip;
					const uint32_t q = fip + lane;
					uint32_t lim = ip - anchor;
					uint32_t bne = 1u;
					bool fne = true;

					if (mp < lim)
						lim = mp;
					if (lane < limit)
						fne = d[fip + lane] != d[fm + lane];
					if (q <= mflimit) {
						uint32_t b4q, b4p;
						const uint32_t sq = lz4e_sld_seq5(da, q + dm, b4q);
						const uint32_t sp = lz4e_sld_seq5(da, q - 2u + dm, b4p);

						q_h = lz4e_hash(sq, b4q, TT);
						q_h2 = lz4e_hash(sp, b4p, TT);
						q_m = q_h2 == q_h ? q - 2u : tab.get(q_h);
						q_ok = (NEAR || q_m + 65535u >= q) && lz4e_sld32u(da, q_m + dm) == sq;
					}
					if (lim) { /* uniform: 0 after an immediate match */
						bool beq = false;

						if (lane < lim)
							beq = d[ip - 1u - lane] == d[mp - 1u - lane];
						bne = lz4e_ballot(LZ4E_FULL, !beq);
						back = bne ? (uint32_t)lz4e_ffs(bne) - 1u : 32u;
					}
					const uint32_t fnz = lz4e_ballot(LZ4E_FULL, fne);

					while (bne == 0 && back < lim) { /* rare: > 32 back */
						const uint32_t i = back + lane;
						bool eq = false;

						if (i < lim)
							eq = d[ip - 1u - i] == d[mp - 1u - i];
						const uint32_t ne = lz4e_ballot(LZ4E_FULL, !eq);

						if (ne) {
							back += (uint32_t)lz4e_ffs(ne) - 1u;
							break;
						}
						back += 32u;
					}
					if (fnz) {
						ml = (uint32_t)lz4e_ffs(fnz) - 1u;
						have = true; /* lane ml holds the next test */
					} else {
						have = false;
						/* >= 32 equal bytes: continue 4 bytes per lane */
						ml = 32u;
						for (;;) {
							const uint32_t o = ml + 4u * lane;
							uint32_t x = 1;

							if (o < limit)
								x = lz4e_sld32u(da, fip + o + dm) ^
								    lz4e_sld32u(da, fm + o + dm);
							const uint32_t nz =
								lz4e_ballot(LZ4E_FULL, x != 0);
							if (nz) {
								const int fl = lz4e_ffs(nz) - 1;
								const uint32_t xf =
									lz4e_shfl(LZ4E_FULL, x, fl);
								ml += 4u * (uint32_t)fl +
								      (((uint32_t)lz4e_ffs(xf) - 1u) >> 3);
								break;
							}
							ml += 128u;
						}
						if (ml > limit)
							ml = limit;
					}
				}

				/* ---- emit one sequence (:352-450) ---- */
				{
					const uint32_t lit = ip - back - anchor;
					const uint32_t mc = back + ml;
					const uint32_t off = ip - mp;
					uint32_t lit_ext = 0, mc_ext = 0;

					if (lit >= 15u)
						lit_ext = (lit - 15u) / 255u + 1u;
					if (mc >= 15u)
						mc_ext = (mc - 15u) / 255u + 1u;
					const uint32_t need = 1u + lit_ext + lit + 2u + mc_ext;

					if (limited &&
					    (op + 1u + lit + 8u + lit / 255u > cap ||
					     op + 1u + lit_ext + lit + 2u + 6u + (mc >> 8) > cap)) {
						/* :358-363 is not on the path of an immediate match
						 * (goto _next_match, :489-492): ip == anchor there */
						const bool first = ip != anchor &&
								   op + 1u + lit + 8u + lit / 255u > cap;

						lz4e_compress_overflow<DIRECT>(st.cur0, st.base, st.win, st.dst.v,
									       st.dst.done, d, op, anchor, lit,
									       lit_ext, off, !first, lane);
						src_end = first ? ip - back : ip + 4u + ml;
						dst_end = first ? op + 1u : op + 1u + lit_ext + lit + 2u;
						return 0;
					}

					bool stream = false;

					if (!whole && op - st.win + need > st.cap) {
						lz4e_stage_flush(st, op, a.out_cap, lane);
						stream = need > st.cap;
					}
					uint8_t *o = st.cur0 + (op - st.win);

					if (!stream && lit < 15u && mc < 15u) {
						/* the common short sequence: token, literals,
						 * offset in one pass, one byte per lane */
						uint8_t v = 0;

						if (lane == 0)
							v = (uint8_t)((lit << 4) | mc);
						else if (lane <= lit)
							v = d[anchor + lane - 1u];
						else
							v = (uint8_t)(lane == lit + 1u ? off : off >> 8);
						if (lane < lit + 3u)
							o[lane] = v;
					} else {
						if (lane == 0)
							o[0] = (uint8_t)(((lit < 15u ? lit : 15u) << 4) |
									 (mc < 15u ? mc : 15u));
						if (lit_ext)
							lz4e_put_run(o + 1, lit_ext, (lit - 15u) % 255u,
								     lane);
						o += 1u + lit_ext;
						if (stream) {
							/* literal run larger than the staging area:
							 * flush the header, stream the literals
							 * straight from the gathered input */
							lz4e_stage_flush(st, op + 1u + lit_ext, a.out_cap,
									 lane);
							lz4e_scatter<32>(st.dst, d + anchor, lit, lane);
							lz4e_stage_rebase(st, op + 1u + lit_ext + lit,
									  a.out_cap);
							o = st.cur0;
						} else {
							if (lit <= 32u) {
								if (lane < lit)
									o[lane] = d[anchor + lane];
							} else {
								lz4e_copy<32>(o, d + anchor, lit, lane);
							}
							o += lit;
						}
						if (lane < 2u)
							o[lane] = (uint8_t)(lane ? off >> 8 : off);
						if (mc_ext)
							lz4e_put_run(o + 2, mc_ext, (mc - 15u) % 255u,
								     lane);
					}
					op += need;
				}

				anchor = ip + 4u + ml;
				if (anchor > mflimit) { /* :456 */
					src_end = anchor;
					break;
				}
				p0 = anchor + 1u;
				if (have) {
					/* the lane that assumed length ml did the work of
					 * :461-493 for this anchor; apply its table writes in
					 * the reference's order */
					const uint32_t okm = lz4e_ballot(LZ4E_FULL, q_ok);

					mp = lz4e_shfl(LZ4E_FULL, q_m, (int)ml);
					if (lane == ml) {
						tab.put(q_h2, anchor - 2u);
						tab.put(q_h, anchor);
					}
					ip = anchor;
					have = (okm >> ml) & 1u; /* else: search from ip + 1 */
					kbase = 0;
				} else {
					kbase = -2;
				}
				lz4e_syncwarp(LZ4E_FULL); /* table writes visible */
			}
		}

		/* ---- last literals (:500-530) ---- */
		{
			const uint32_t run = n - anchor;
			const uint32_t ext = run >= 15u ? (run - 15u) / 255u + 1u : 0u;

			if (limited && op + run + 1u + (run + 240u) / 255u > cap) {
				/* :505-509: everything before the last run is out */
				lz4e_compress_overflow<DIRECT>(st.cur0, st.base, st.win, st.dst.v, st.dst.done,
						       d, op, anchor, 0, 0, 0, false, lane);
				dst_end = op;
				return 0;
			}
			if (!DIRECT && op - st.win + 1u + ext > st.cap)
				lz4e_stage_flush(st, op, a.out_cap, lane);
			uint8_t *o = st.cur0 + (op - st.win);

			if (lane == 0)
				o[0] = (uint8_t)((run < 15u ? run : 15u) << 4);
			lz4e_put_run(o + 1, ext, (run - 15u) % 255u, lane);
			op += 1u + ext;
			dst_end = op;
			if (DIRECT) {
				lz4e_copy<32>(o + 1u + ext, d + anchor, run, lane);
				op += run;
			} else if (op - st.win + run > st.cap) {
				lz4e_stage_flush(st, op, a.out_cap, lane);
				lz4e_scatter<32>(st.dst, d + anchor, run, lane);
				op += run;
			} else {
				lz4e_copy<32>(o + 1u + ext, d + anchor, run, lane);
				op += run;
				lz4e_syncwarp(LZ4E_FULL);
				lz4e_scatter<32>(st.dst, st.cur0, op - st.win, lane);
			}
			ret = (int32_t)op;
		}
	}

	return ret;
}

/*
 * Compress block b.  sm points at this warp's shared memory
 * (lz4e_compress_smem_bytes()).  All 32 lanes call it together.
 */
template <bool DIRECT, bool GIN, bool TMA>
LZ4E_DEV void lz4e_compress_block(const lz4e_compress_args &a, uint32_t b,
				  uint8_t *sm, uint32_t lane, uint32_t slot,
				  uint64_t *bar, uint32_t &phase)
{
	const struct lz4e_b200_dev_iter si = a.src_iters[b];
	const struct lz4e_b200_dev_iter di = a.dst_iters[b];
	const uint32_t n = si.bi_size;
	uint8_t *in_area = GIN ? a.in_scratch + (size_t)slot * a.in_stride
			       : sm + a.table_bytes;
	uint8_t *out_area = sm + a.table_bytes + (GIN ? 0u : a.in_cap + LZ4E_K_PAD);
	const uint8_t *d; /* d[pos] = source byte at linear position pos */
	int tt = LZ4E_TT_U16;
	int32_t ret = 0;
	uint32_t src_end = 0, dst_end = 0;
	bool p12 = false;

	if (n > LZ4E_K_MAX_INPUT) /* lz4e_compress.c:245-248 */
		goto done;
	/* the staging window must hold the longest length run of a block by itself
	 * (token or offset plus one length byte per 255 bytes, lz4e_compress.c:
	 * 365-377, :432-446, :511-522): the launch sizes out_cap from in_cap */
	if (n > a.in_cap || (!DIRECT && n / 255u + 20u > a.out_cap)) {
		if (lane == 0)
			lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
		goto done;
	}
	if (DIRECT && di.bi_size) {
		/* promise of LZ4E_B200_DEV_DST_LINEAR: one segment holds the whole
		 * capacity */
		const struct lz4e_b200_dev_bvec *dv = a.dst_bvecs + di.first + di.bi_idx;

		if (dv->len < di.bi_bvec_done || dv->len - di.bi_bvec_done < di.bi_size) {
			if (lane == 0)
				lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
			goto done;
		}
	}

	/*
	 * ---- gather + address mode (lz4e_compress.c:184-211) ----
	 * Shared-memory kernels: every segment whose address is congruent mod 16
	 * with its place in the staging area goes through the SM's TMA unit (1-D
	 * cp.async.bulk, 16-byte aligned body; head and tail bytes by the lanes);
	 * the copies are issued by lane 0, the warp zeroes its hash table while
	 * they are in flight and then waits on the mbarrier.  Segments that are
	 * not congruent (odd offsets behind the first one) are moved by the lanes
	 * with funnel-shifted words as before.
	 */
	{
		const struct lz4e_b200_dev_bvec *v =
			a.src_bvecs + si.first + si.bi_idx;
		uint32_t done_in = si.bi_bvec_done, rem = n, pos = 0, i = 0;
		bool too_many = false;
		bool bulk = false;

		(void)bulk;
		uint8_t *dbase = in_area;

		if (n)
			dbase += (uint32_t)(v->addr + done_in) & 15u;
		d = dbase;
#if defined(LZ4E_HAVE_TMA)
		if (TMA && lane == 0)
			lz4e_fence_async_smem(); /* the previous block's reads are done */
#endif
		while (rem) {
			if (i >= BIO_MAX_VECS) {
				too_many = true;
				break;
			}
			const uint32_t len = v[i].len;
			uint32_t take = len - done_in;
			uint8_t *g = (uint8_t *)(uintptr_t)(v[i].addr + done_in);

			LZ4E_ASSUME_GLOBAL(g);
			if (i >= 16 || len > 4096u)
				tt |= LZ4E_TT_U32;
			if (len > (1u << 24))
				tt |= LZ4E_TT_U64;
			if (take > rem)
				take = rem;
#if defined(LZ4E_HAVE_TMA)
			if (TMA && take >= 256u &&
			    ((((uint32_t)(uintptr_t)g) ^ lz4e_smem_addr(dbase + pos)) & 15u) == 0) {
				const uint32_t head = (uint32_t)(-(uintptr_t)g) & 15u;
				const uint32_t body = (take - head) & ~15u;
				const uint32_t tail = take - head - body;
				uint8_t *o = dbase + pos;

				if (lane == 0) {
					lz4e_mbar_expect(bar, body);
					lz4e_tma_load(o + head, g + head, body, bar);
				}
				if (lane < head)
					o[lane] = g[lane];
				if (lane < tail)
					o[head + body + lane] = g[head + body + lane];
				bulk = true;
			} else
#endif
				lz4e_copy<32>(dbase + pos, g, take, lane);
			pos += take;
			rem -= take;
			done_in = 0;
			i++;
		}
#if defined(LZ4E_HAVE_TMA)
		if (TMA && bulk && lane == 0)
			lz4e_mbar_arrive(bar);
#endif
		/* packed table: the launch reserved 12 KiB, this block is byU16 */
		p12 = !GIN && tt == LZ4E_TT_U16 && a.table_bytes == LZ4E_K_P12_BYTES && n <= 4096u;
		if (!(too_many && n >= 13) && n >= 13) {
			/* zero the table while the bulk copies land */
			const uint32_t tbytes =
				p12 ? LZ4E_K_P12_BYTES
				    : (tt == LZ4E_TT_U16 ? 8192u : tt == LZ4E_TT_U32 ? 4096u : 2048u) *
					      (n > 65536u ? 4u : 2u);
			const lz4e_u128 z = { 0, 0, 0, 0 };

			if (tbytes <= a.table_bytes)
				for (uint32_t k = lane * 16u; k < tbytes; k += 512u)
					lz4e_st128(sm + k, z);
		}
#if defined(LZ4E_HAVE_TMA)
		if (TMA && bulk) {
			lz4e_mbar_wait(bar, phase);
			phase ^= 1u;
		}
#endif
		if (too_many && n >= 13) /* :268-277 */
			goto done;
		lz4e_syncwarp(LZ4E_FULL); /* gathered bytes and table visible to all lanes */
	}

	if (n > 65536u) {
		/* positions no longer fit 16 bits: same table sizes in entries
		 * (lz4e_compress.c:48-57), 32-bit linear positions (the launch
		 * reserved 16 KiB of table for them).  byU16 cannot
		 * occur here: it needs <= 16 bvecs of <= 4096 bytes (:200-202) */
		if (tt == LZ4E_TT_U32)
			ret = lz4e_compress_encode<LZ4E_TT_U32, DIRECT, GIN, uint32_t>(
				a, di, d, n, sm, out_area, lane, src_end, dst_end);
		else if (tt == LZ4E_TT_U64)
			ret = lz4e_compress_encode<LZ4E_TT_U64, DIRECT, GIN, uint32_t>(
				a, di, d, n, sm, out_area, lane, src_end, dst_end);
		else if (lane == 0)
			lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
		goto done;
	}
	switch (tt) {
	case LZ4E_TT_U16:
		if (!GIN && p12)
			ret = lz4e_compress_encode<LZ4E_TT_U16, DIRECT, GIN, uint16_t, true>(
				a, di, d, n, sm, out_area, lane, src_end, dst_end);
		else
			ret = lz4e_compress_encode<LZ4E_TT_U16, DIRECT, GIN, uint16_t>(
				a, di, d, n, sm, out_area, lane, src_end, dst_end);
		break;
	case LZ4E_TT_U32:
		ret = lz4e_compress_encode<LZ4E_TT_U32, DIRECT, GIN, uint16_t>(
			a, di, d, n, sm, out_area, lane, src_end, dst_end);
		break;
	default:
		ret = lz4e_compress_encode<LZ4E_TT_U64, DIRECT, GIN, uint16_t>(
			a, di, d, n, sm, out_area, lane, src_end, dst_end);
		break;
	}

done:
	if (lane == 0) {
		a.rets[b] = ret;
		if (a.ends) {
			/* also after a 0 return: the reference leaves its iterators
			 * where the failing check found them */
			a.ends[b].src_pos = src_end;
			a.ends[b].dst_pos = dst_end;
		}
	}
	/* the next block reuses this warp's shared memory */
	lz4e_syncwarp(LZ4E_FULL);
}

/* persistent loop: warps pull block indices from a global counter */
template <bool DIRECT, bool GIN, bool TMA = false>
LZ4E_DEV void lz4e_compress_warp(const lz4e_compress_args &a, uint8_t *sm,
				 uint32_t lane, uint32_t slot)
{
	/* the warp's mbarrier sits at the end of its shared-memory area */
	uint64_t *bar = reinterpret_cast<uint64_t *>(sm + a.warp_smem - LZ4E_K_BAR);
	uint32_t phase = 0;

#if defined(LZ4E_HAVE_TMA)
	if (TMA && lane == 0)
		lz4e_mbar_init(bar, 1);
	lz4e_syncwarp(LZ4E_FULL);
#endif
	for (;;) {
		uint32_t b = 0;

		if (lane == 0)
			b = lz4e_atomic_add(a.work_counter, 1u);
		b = lz4e_shfl(LZ4E_FULL, b, 0);
		if (b >= a.nblocks)
			break;
		if (a.order)
			b = a.order[b];
		lz4e_compress_block<DIRECT, GIN, TMA>(a, b, sm, lane, slot, bar, phase);
	}
}

#endif /* LZ4E_COMPRESS_BODY_CUH */
