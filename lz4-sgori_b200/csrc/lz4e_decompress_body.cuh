/*
 * lz4e_decompress_body.cuh — a group of GS lanes (GS = 8, 16 or 32) decodes
 * one LZ4 block; 32/GS blocks are in flight per warp.
 *
 * Semantics: LZ4E_decompress_safe (lz4e/lz4e_decompress.c:462-470 of the
 * reference), i.e. LZ4E_decompress_generic (:62-460) with endOnInputSize,
 * decode_full_block, noDict, lowPrefix == dest — including which malformed
 * inputs are rejected and the value returned for them, -(input offset) - 1
 * (:458-459).  The reference function takes contiguous buffers; here source and
 * destination are scatter-gather lists (the SG decoder north_star asks for).
 *
 * Control flow is uniform inside a group: every lane parses the same token,
 * lengths and offset (broadcast loads), then lanes split the literal and match
 * bytes between them.  The two-stage "shortcut" of :150-191 is kept only as
 * the predicate that decides which end-of-block checks apply, because that is
 * observable in the return value for malformed streams.
 *
 * Overlapping matches (offset < match length) are resolved arithmetically:
 * out[op + i] = out[op - offset + i % offset] reads only bytes that were final
 * before the match started, so all lanes can copy at once (the
 * replicated-pattern path).  offset == 0 yields zero bytes, exactly what
 * LZ4_write32(op, offset) at :309-314 makes the reference produce.
 *
 * STAGE = true : the block's output is built in shared memory and written to
 *                the destination segments once, coalesced (blocks up to
 *                stage_cap bytes, e.g. the 4 KiB chunk size of lz4e_bdev).
 * STAGE = false: output goes straight to global memory; if the destination is
 *                one contiguous segment it is used in place, otherwise the
 *                block is decoded into a per-group scratch slot and scattered.
 */
#ifndef LZ4E_DECOMPRESS_BODY_CUH
#define LZ4E_DECOMPRESS_BODY_CUH

#include "../../include/lz4e_b200.h"
#include "lz4e_copy.cuh"
#include "lz4e_simt.h"

#define LZ4E_D_PAD 32u
#ifndef LZ4E_FAULT_RANGE
#define LZ4E_FAULT_RANGE 1u
#endif

struct lz4e_decompress_args {
	const struct lz4e_b200_dev_bvec *src_bvecs;
	const struct lz4e_b200_dev_iter *src_iters;
	const struct lz4e_b200_dev_bvec *dst_bvecs;
	const struct lz4e_b200_dev_iter *dst_iters;
	int32_t *rets;
	uint32_t *work_counter;
	uint32_t *faults;
	uint32_t nblocks;
	uint32_t stage_cap;	 /* STAGE: shared-memory bytes per group */
	uint8_t *scratch;	 /* !STAGE: linear slots for non-contiguous dst */
	uint32_t scratch_stride; /* bytes per slot */
	/* lane-per-block decoder only (lz4e_decompress_lane.cuh): linear slots
	 * for compressed blocks whose source is not one contiguous segment */
	uint8_t *in_scratch;
	uint32_t in_stride;
	/* work item i is block order[i] (cost-ordered queue, lz4e_order.cu);
	 * NULL = batch order */
	const uint32_t *order;
	/* wave decoder: waves whose sequences average at least this many output
	 * bytes are copied one sequence at a time (0: never) */
	uint32_t wave_serial;
};


/* forward-only view of the compressed SG stream: byte p lives at bias[p]
 * while lo <= p < hi */
struct lz4e_in_view {
	const uint8_t *bias;
	const struct lz4e_b200_dev_bvec *v;
	uint32_t hi;
	uint32_t end; /* compressed size */
};

LZ4E_DEV void lz4e_in_open(lz4e_in_view &in,
			   const struct lz4e_b200_dev_bvec *table,
			   const struct lz4e_b200_dev_iter &it)
{
	in.v = table + it.first + it.bi_idx;
	in.end = it.bi_size;
	if (it.bi_size == 0) {
		in.hi = 0;
		in.bias = 0;
		return;
	}
	const uint32_t room = in.v->len - it.bi_bvec_done;

	in.hi = room < it.bi_size ? room : it.bi_size;
	in.bias = (const uint8_t *)(uintptr_t)(in.v->addr + it.bi_bvec_done);
	LZ4E_ASSUME_GLOBAL(in.bias);
}

/* make the segment holding position p current (p < end) */
LZ4E_DEV void lz4e_in_seek(lz4e_in_view &in, uint32_t p)
{
	while (p >= in.hi && in.hi < in.end) {
		const uint32_t lo = in.hi;

		in.v++;
		in.hi = lo + in.v->len;
		if (in.hi > in.end || in.hi < lo)
			in.hi = in.end;
		in.bias = (const uint8_t *)(uintptr_t)(in.v->addr - lo);
		LZ4E_ASSUME_GLOBAL(in.bias);
	}
}

LZ4E_DEV uint32_t lz4e_in_rd8_slow(const lz4e_in_view &in, uint32_t p)
{
	const struct lz4e_b200_dev_bvec *w = in.v + 1;
	uint32_t base = in.hi;

	for (;;) {
		if (p - base < w->len) {
			const uint8_t *g =
				(const uint8_t *)(uintptr_t)(w->addr + (p - base));
			LZ4E_ASSUME_GLOBAL(g);
			return *g;
		}
		base += w->len;
		w++;
	}
}

LZ4E_DEV uint32_t lz4e_in_rd8(const lz4e_in_view &in, uint32_t p)
{
	if (p < in.hi)
		return in.bias[p];
	return lz4e_in_rd8_slow(in, p);
}

/* literal run: `len` bytes from compressed position ip to out (W lanes) */
template <int W>
LZ4E_DEV void lz4e_dgroup_literals(const lz4e_in_view &in, uint8_t *out,
				   uint32_t ip, uint32_t len, uint32_t l)
{
	if (len >= 32u && ip + len <= in.hi) {
		lz4e_copy<W>(out, in.bias + ip, len, l);
		return;
	}
	if (len < 32u) {
		for (uint32_t i = l; i < len; i += W)
			out[i] = (uint8_t)lz4e_in_rd8(in, ip + i);
		return;
	}
	/* a long run over several source segments: piece by piece */
	lz4e_in_view w = in;
	uint32_t done = 0;

	while (done < len) {
		const uint32_t p = ip + done;

		lz4e_in_seek(w, p);
		uint32_t take = w.hi - p;

		if (take > len - done)
			take = len - done;
		lz4e_copy<W>(out + done, w.bias + p, take, l);
		done += take;
	}
}

#define LZ4E_P_LIT 1u
#define LZ4E_P_MATCH 2u
#define LZ4E_P_FINISH 4u
#define LZ4E_D_WIDE 128u /* copies of at least this many bytes go to the warp */
#ifndef LZ4E_D_SERVICE
#define LZ4E_D_SERVICE 0 /* 1: long copies are served by all 32 lanes of the warp */
#endif

LZ4E_DEV uint64_t lz4e_shfl64(uint64_t v, int src)
{
	const uint32_t lo = lz4e_shfl(LZ4E_FULL, (uint32_t)v, src);
	const uint32_t hi = lz4e_shfl(LZ4E_FULL, (uint32_t)(v >> 32), src);

	return ((uint64_t)hi << 32) | lo;
}

/* match copy by all 32 lanes: o[i] = o[i - dist]; dist == 0 writes zeros
 * (lz4e_decompress.c:309-314).  An overlapping match is a pattern of period
 * dist: whole periods are written byte-wise until >= 64 bytes exist, then the
 * finished part is doubled with vector copies. */
LZ4E_DEV void lz4e_wide_match(uint8_t *o, uint32_t dist, uint32_t len, uint32_t lane)
{
	if (dist == 0u) {
		for (uint32_t i = lane; i < len; i += 32u)
			o[i] = 0;
		return;
	}
	if (dist >= len) {
		lz4e_copy<32>(o, o - dist, len, lane);
		return;
	}
	const uint8_t *ms = o - dist;
	uint32_t have = dist * ((63u + dist) / dist);

	if (have > len)
		have = len;
	for (uint32_t i = lane; i < have; i += 32u)
		o[i] = ms[i % dist];
	while (have < len) {
		uint32_t chunk = len - have;

		if (chunk > have)
			chunk = have;
		lz4e_syncwarp(LZ4E_FULL);
		lz4e_copy<32>(o + have, o, chunk, lane);
		have += chunk;
	}
}

template <int GS, bool STAGE>
struct lz4e_dgroup {
	lz4e_in_view in;
	uint8_t *out;	/* out[op]: linear output (shared or global) */
	uint32_t b;	/* block index */
	int32_t iend;	/* compressed size */
	int32_t oend;	/* output capacity */
	uint32_t ip, op;
	bool active;	/* a block is being decoded */
	bool done;	/* the batch has no more blocks for this group */
	/* long copies handed to the whole warp (see lz4e_dwarp_service) */
	uint32_t pend;	/* LZ4E_P_LIT | LZ4E_P_MATCH | LZ4E_P_FINISH */
	const uint8_t *p_lsrc;
	uint32_t p_lop, p_llen;		/* literals: out + p_lop <- p_lsrc */
	uint32_t p_mop, p_mdist, p_mlen; /* match: out + p_mop, distance, length */
	int32_t p_ret;
	bool scatter;	/* !STAGE: output sits in scratch, scatter when done */
};

/* write the result of a finished block and move its bytes home */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_dgroup_finish(const lz4e_decompress_args &a,
				 lz4e_dgroup<GS, STAGE> &g, int32_t ret,
				 uint32_t gmask, uint32_t gl)
{
	lz4e_syncwarp(gmask);
	if ((STAGE || g.scatter) && ret > 0) {
		lz4e_cursor c = lz4e_cursor_init(a.dst_bvecs, a.dst_iters[g.b]);

		lz4e_scatter<GS>(c, g.out, (uint32_t)ret, gl);
	}
	if (gl == 0)
		a.rets[g.b] = ret;
	g.active = false;
	lz4e_syncwarp(gmask); /* staging is reused by the next block */
}

/*
 * sm: this warp's shared memory ((32/GS) * (stage_cap + LZ4E_D_PAD) bytes when
 * STAGE, unused otherwise); slot0: index of this warp's first scratch slot.
 */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_decompress_warp_impl(const lz4e_decompress_args &a,
					uint8_t *sm, uint32_t slot0,
					uint32_t lane)
{
	const uint32_t gl = lane & (GS - 1);  /* lane inside the group */
	const uint32_t gi = lane / GS;	      /* group inside the warp */
	const uint32_t gmask =
		(GS == 32 ? 0xffffffffu : ((1u << GS) - 1u)) << (gi * GS);
	const int gleader = (int)(gi * GS);
	lz4e_dgroup<GS, STAGE> g;

	g.active = false;
	g.scatter = false;
	g.done = false;
	g.pend = 0;
	g.p_lsrc = 0;
	g.p_lop = g.p_llen = g.p_mop = g.p_mdist = g.p_mlen = 0;
	g.p_ret = 0;
	for (;;) {
		/*
		 * Every iteration starts with a full-warp vote.  It ends the loop
		 * when all groups have run out of blocks and, just as important,
		 * it makes the 32/GS groups of the warp RECONVERGE once per
		 * sequence: they then issue the (branch-free) fast path together,
		 * one instruction stream for 32/GS blocks.  Groups that are done
		 * keep voting until the others finish.
		 */
		if (lz4e_ballot(LZ4E_FULL, !g.done) == 0)
			break;
		/*
		 * Long literal runs and long matches are not copied by the 8/16
		 * lanes of their group: the group posts them and ALL 32 lanes of
		 * the warp (groups that are idle included) move them here with
		 * 16-byte vectors.
		 */
		if (LZ4E_D_SERVICE) {
			uint32_t pm = lz4e_ballot(LZ4E_FULL, g.pend != 0u && gl == 0u);

			if (pm) {
				lz4e_syncwarp(LZ4E_FULL); /* group-written bytes visible */
				do {
					const int L = lz4e_ffs(pm) - 1;
					uint8_t *out = (uint8_t *)(uintptr_t)lz4e_shfl64(
						(uint64_t)(uintptr_t)g.out, L);
					const uint8_t *lsrc = (const uint8_t *)(uintptr_t)lz4e_shfl64(
						(uint64_t)(uintptr_t)g.p_lsrc, L);
					const uint32_t what = lz4e_shfl(LZ4E_FULL, g.pend, L);
					const uint32_t lop = lz4e_shfl(LZ4E_FULL, g.p_lop, L);
					const uint32_t llen = lz4e_shfl(LZ4E_FULL, g.p_llen, L);
					const uint32_t mop = lz4e_shfl(LZ4E_FULL, g.p_mop, L);
					const uint32_t mdist = lz4e_shfl(LZ4E_FULL, g.p_mdist, L);
					const uint32_t mlen = lz4e_shfl(LZ4E_FULL, g.p_mlen, L);

					if (!STAGE)
						LZ4E_ASSUME_GLOBAL(out);
					if (what & LZ4E_P_LIT) {
						LZ4E_ASSUME_GLOBAL(lsrc);
						lz4e_copy<32>(out + lop, lsrc, llen, lane);
					}
					if (what & LZ4E_P_MATCH) {
						lz4e_syncwarp(LZ4E_FULL);
						lz4e_wide_match(out + mop, mdist, mlen, lane);
					}
					pm &= pm - 1u;
				} while (pm);
				lz4e_syncwarp(LZ4E_FULL);
			}
			if (g.pend) {
				const uint32_t what = g.pend;

				g.pend = 0;
				if (what & LZ4E_P_FINISH)
					lz4e_dgroup_finish(a, g, g.p_ret, gmask, gl);
			}
		}
		/* what this group asks the uniform copy phase below to do */
		uint32_t c_lit = 0, c_mlen = 0, c_off = 1;
		const uint8_t *c_lsrc = nullptr;
		uint8_t *c_o = nullptr;

		do { /* one group step; `break` leaves it */
		if (g.done)
			break;
		if (!g.active) {
			/* ---- next block for this group ---- */
			uint32_t b = 0;

			if (gl == 0)
				b = lz4e_atomic_add(a.work_counter, 1u);
			b = lz4e_shfl(gmask, b, gleader);
			if (b >= a.nblocks) {
				g.done = true;
				break;
			}
			const struct lz4e_b200_dev_iter si = a.src_iters[b];
			const struct lz4e_b200_dev_iter di = a.dst_iters[b];

			g.b = b;
			g.iend = (int32_t)si.bi_size;
			g.oend = (int32_t)di.bi_size;
			g.ip = g.op = 0;
			g.active = true;
			g.scatter = false;
			if (si.bi_size > (16u << 20) || di.bi_size > 0x7fffffffu ||
			    (STAGE && di.bi_size > a.stage_cap)) {
				if (gl == 0) {
					lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
					a.rets[b] = -1;
				}
				g.active = false;
				break;
			}
			/* special cases, lz4e_decompress.c:113-120 */
			if (g.oend == 0 || g.iend == 0) {
				int32_t r = -1;

				lz4e_in_open(g.in, a.src_bvecs, si);
				if (g.oend == 0 && g.iend == 1 &&
				    lz4e_in_rd8(g.in, 0) == 0)
					r = 0;
				if (gl == 0)
					a.rets[b] = r;
				g.active = false;
				break;
			}
			lz4e_in_open(g.in, a.src_bvecs, si);
			const struct lz4e_b200_dev_bvec *dv =
				a.dst_bvecs + di.first + di.bi_idx;

			if (STAGE) {
				/* congruent mod 16 with the first destination
				 * segment, so the final scatter is vector stores */
				g.out = sm + gi * (a.stage_cap + LZ4E_D_PAD) +
					((uint32_t)(dv->addr + di.bi_bvec_done) & 15u);
			} else {

				if (dv->len - di.bi_bvec_done >= di.bi_size) {
					g.out = (uint8_t *)(uintptr_t)(dv->addr +
								       di.bi_bvec_done);
				} else {
					if (di.bi_size > a.scratch_stride) {
						if (gl == 0) {
							lz4e_atomic_or(a.faults,
								       LZ4E_FAULT_RANGE);
							a.rets[b] = -1;
						}
						g.active = false;
						break;
					}
					g.out = a.scratch + (size_t)(slot0 + gi) *
								    a.scratch_stride;
					g.scatter = true;
				}
				LZ4E_ASSUME_GLOBAL(g.out);
			}
		}

		/* ---- one sequence ---- */
		const int32_t iend = g.iend, oend = g.oend;
		uint32_t ip = g.ip, op = g.op;

		lz4e_in_seek(g.in, ip);

		/*
		 * Common path: a sequence with at most one extension byte per
		 * length (literals < 270, match < 274 bytes) that lies inside the
		 * current source segment and far enough from both ends that none of
		 * the reference's end-of-block rules can fire:
		 *   next ip <= iend - 12  rules out the input checks at
		 *                         lz4e_decompress.c:197-200, :226-227, :322
		 *   op + lit + mlen <= oend - 12  rules out :226, :425-431
		 * so the shortcut predicate of :150-156 makes no difference here
		 * and the sequence is just "copy literals, copy match".  All groups
		 * of the warp run this path together; anything else (long runs,
		 * block tail, segment crossings, malformed input) takes the general
		 * path below, which applies the rules one by one.
		 */
		{
			const int32_t hi = (int32_t)g.in.hi < iend ? (int32_t)g.in.hi : iend;
			const uint8_t *inb = g.in.bias;

			if ((int32_t)ip + 4 <= hi) {
				const uint32_t tok = inb[ip];
				const uint32_t e1 = inb[ip + 1u];
				const uint32_t lext = (tok >> 4) == 15u;
				const uint32_t lit = (tok >> 4) + (lext ? e1 : 0u);
				const uint32_t lp = ip + 1u + lext; /* first literal */
				const uint32_t fp = lp + lit;	     /* offset field */

				if (!(lext && e1 == 255u) && (int32_t)fp + 15 <= hi) {
					const uint32_t off = (uint32_t)inb[fp] |
							     ((uint32_t)inb[fp + 1u] << 8);
					const uint32_t e2 = inb[fp + 2u];
					const uint32_t mext = (tok & 15u) == 15u;
					const uint32_t mlen = (tok & 15u) + 4u + (mext ? e2 : 0u);
					const uint32_t nip = fp + 2u + mext;

					if (!(mext && e2 == 255u) && off != 0u && off <= op + lit &&
					    lit < LZ4E_D_WIDE && mlen < LZ4E_D_WIDE &&
					    (int64_t)op + lit + mlen <= (int64_t)oend - 12) {
						c_o = g.out + op;
						c_lsrc = inb + lp;
						c_lit = lit;
						c_off = off;
						c_mlen = mlen;
						g.ip = nip;
						g.op = op + lit + mlen;
						break;
					}
				}
			}
		}

		const uint32_t token = lz4e_in_rd8(g.in, ip++);
		uint32_t len = token >> 4;
		const bool shortcut = len != 15u && (int32_t)ip < iend - 16 &&
				      (int32_t)op <= oend - 32; /* :150-156 */

		if (!shortcut) {
			if (len == 15u) { /* :194-220 */
				uint32_t s;

				if ((int32_t)ip >= iend - 15) {
					lz4e_dgroup_finish(a, g, -(int32_t)ip - 1, gmask, gl);
					break;
				}
				do {
					s = lz4e_in_rd8(g.in, ip++);
					len += s;
				} while ((int32_t)ip < iend - 15 && s == 255u);
			}
			/* :222-266 */
			if ((int64_t)op + len > (int64_t)oend - 12 ||
			    (int64_t)ip + len > (int64_t)iend - 8) {
				if ((int64_t)ip + len != (int64_t)iend ||
				    (int64_t)op + len > (int64_t)oend) {
					lz4e_dgroup_finish(a, g, -(int32_t)ip - 1, gmask, gl);
					break;
				}
				if (LZ4E_D_SERVICE && len >= LZ4E_D_WIDE && ip + len <= g.in.hi) {
					/* the whole warp copies, then the block ends */
					g.pend = LZ4E_P_LIT | LZ4E_P_FINISH;
					g.p_lsrc = g.in.bias + ip;
					g.p_lop = op;
					g.p_llen = len;
					g.p_ret = (int32_t)(op + len);
					break;
				}
				lz4e_dgroup_literals<GS>(g.in, g.out + op, ip, len, gl);
				lz4e_dgroup_finish(a, g, (int32_t)(op + len), gmask, gl);
				break;
			}
		}
		const bool wide_lit = LZ4E_D_SERVICE && len >= LZ4E_D_WIDE && ip + len <= g.in.hi;

		if (wide_lit) {
			g.p_lsrc = g.in.bias + ip;
			g.p_lop = op;
			g.p_llen = len;
		} else {
			lz4e_dgroup_literals<GS>(g.in, g.out + op, ip, len, gl);
		}
		ip += len;
		op += len;

		/* offset, :280-283 */
		const uint32_t offset =
			lz4e_in_rd8(g.in, ip) | (lz4e_in_rd8(g.in, ip + 1) << 8);
		uint32_t mlen = token & 15u;

		ip += 2;
		if (offset > op) { /* :299-302 */
			lz4e_dgroup_finish(a, g, -(int32_t)ip - 1, gmask, gl);
			break;
		}
		const bool full_shortcut = shortcut && mlen != 15u && offset >= 8u;

		if (mlen == 15u) { /* :316-334 */
			uint32_t s;
			bool bad = false;

			do {
				s = lz4e_in_rd8(g.in, ip++);
				if ((int32_t)ip > iend - 5) {
					bad = true;
					break;
				}
				mlen += s;
			} while (s == 255u);
			if (bad) {
				lz4e_dgroup_finish(a, g, -(int32_t)ip - 1, gmask, gl);
				break;
			}
		}
		mlen += 4u;
		if (!full_shortcut &&
		    (int64_t)op + mlen > (int64_t)oend - 5) { /* :420-431 */
			lz4e_dgroup_finish(a, g, -(int32_t)ip - 1, gmask, gl);
			break;
		}

		if (wide_lit || (LZ4E_D_SERVICE && mlen >= LZ4E_D_WIDE)) {
			/* literals first, then the match, by all 32 lanes */
			g.pend = (wide_lit ? LZ4E_P_LIT : 0u) | LZ4E_P_MATCH;
			g.p_mop = op;
			g.p_mdist = offset;
			g.p_mlen = mlen;
			g.ip = ip;
			g.op = op + mlen;
			break;
		}
		/* match: literals (and everything earlier) must be visible */
		lz4e_syncwarp(gmask);
		{
			uint8_t *o = g.out + op;

			if (offset == 0u) {
				for (uint32_t i = gl; i < mlen; i += GS)
					o[i] = 0;
			} else if (offset >= mlen) {
				if (mlen >= 64u)
					lz4e_copy<GS>(o, o - offset, mlen, gl);
				else
					for (uint32_t i = gl; i < mlen; i += GS)
						o[i] = (o - offset)[i];
			} else if (mlen < 96u) {
				const uint8_t *ms = o - offset;

				for (uint32_t i = gl; i < mlen; i += GS)
					o[i] = ms[i % offset];
			} else {
				/* long overlapping match = a pattern of period `offset`:
				 * write whole periods byte-wise until at least 32 bytes
				 * exist, then keep doubling with vector copies */
				const uint8_t *ms = o - offset;
				uint32_t have = offset * ((31u + offset) / offset);

				for (uint32_t i = gl; i < have; i += GS)
					o[i] = ms[i % offset];
				while (have < mlen) {
					uint32_t chunk = mlen - have;

					if (chunk > have)
						chunk = have;
					lz4e_syncwarp(gmask);
					lz4e_copy<GS>(o + have, o, chunk, gl);
					have += chunk;
				}
			}
		}
		op += mlen;
		g.ip = ip;
		g.op = op;
		} while (0);

		/*
		 * ---- uniform copy phase: executed by all 32 lanes together ----
		 * Literals, then the match in steps of GS bytes.  The barriers are
		 * full-warp ones (one cheap instruction; a per-group mask would
		 * cost a MATCH/vote sequence each time), so every lane runs the
		 * same number of steps: the maximum over the warp's groups.
		 * With offset >= GS (or no overlap) each step reads bytes that
		 * were final before it started; a shorter offset is a pattern of
		 * period c_off, read at (i mod c_off) with the index kept
		 * incrementally.
		 */
		for (uint32_t i = gl; i < c_lit; i += GS)
			c_o[i] = c_lsrc[i];
		{
			uint8_t *o = c_o + c_lit;
			const uint8_t *ms = o - c_off;

			if (lz4e_ballot(LZ4E_FULL, c_off < c_mlen) == 0) {
				/* no group has an overlapping match: one barrier (the
				 * literals just written may be the source), then every
				 * lane copies its bytes independently */
				lz4e_syncwarp(LZ4E_FULL);
				for (uint32_t i = gl; i < c_mlen; i += GS)
					o[i] = ms[i];
			} else {
				const uint32_t steps = (c_mlen + (uint32_t)GS - 1u) / (uint32_t)GS;
				const uint32_t maxsteps = lz4e_reduce_max(LZ4E_FULL, steps);
				uint32_t j = gl, dj = GS, wrap = 0xffffffffu;

				if (c_off < (uint32_t)GS && c_off < c_mlen) {
					j = gl % c_off;
					dj = (uint32_t)GS % c_off;
					wrap = c_off;
				}
				for (uint32_t st = 0; st < maxsteps; st++) {
					const uint32_t i = st * (uint32_t)GS + gl;

					lz4e_syncwarp(LZ4E_FULL);
					if (i < c_mlen)
						o[i] = ms[j];
					j += dj;
					if (j >= wrap)
						j -= wrap;
				}
			}
		}
	}
}

template <int GS>
LZ4E_DEV void lz4e_decompress_warp(const lz4e_decompress_args &a, uint8_t *sm,
				   uint32_t slot0, uint32_t lane)
{
	if (a.stage_cap)
		lz4e_decompress_warp_impl<GS, true>(a, sm, slot0, lane);
	else
		lz4e_decompress_warp_impl<GS, false>(a, sm, slot0, lane);
}

#endif /* LZ4E_DECOMPRESS_BODY_CUH */
