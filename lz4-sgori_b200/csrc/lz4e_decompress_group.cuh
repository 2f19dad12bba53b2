/*
 * lz4e_decompress_group.cuh — the sub-warp GROUP decoder: GS lanes (4, 8 or 16)
 * decode one LZ4 block, 32/GS blocks are in flight per warp and advance in
 * lockstep, one sequence per step.
 *
 * Semantics: LZ4E_decompress_safe (lz4e/lz4e_decompress.c:462-470 of the
 * reference), i.e. LZ4E_decompress_generic (:62-460) with endOnInputSize,
 * decode_full_block, noDict, lowPrefix == dest — the same accepted streams, the
 * same output and the same -(input offset) - 1 for malformed ones (:458-459).
 * The reference takes contiguous buffers; source and destination here are
 * scatter-gather segment lists.
 *
 * What is different from a transcription of the reference loop:
 *
 *  parse   One unaligned 8-byte window (three aligned word loads + two funnel
 *          shifts) holds token, up to 4 literals, the offset and the first
 *          match-length byte of a common sequence: the chain from one token to
 *          the next is one load and a handful of integer instructions.
 *  output  Scattered destinations are written IN PLACE.  A group keeps the
 *          linear range and address bias of the segment its write cursor is in
 *          and of the one before it in registers, and a small shared-memory
 *          map (segment starts + biases) of the whole block for match sources
 *          further back; nothing is decoded into scratch and copied again.
 *          Blocks of the chunk size (STAGE) are built in shared memory and
 *          leave as 16-byte vector stores.
 *  copies  Literals of the common sequence come out of the parse window
 *          (registers).  A short match is split over the group's lanes, two
 *          bytes per lane; its loads are issued BEFORE the stores of the
 *          previous sequence's match (PIPE), whose data has been in flight
 *          for a whole step, unless the new match reads what those stores
 *          write (then they go first).  Longer or overlapping matches are
 *          copied by the group with vector/word accesses or as a replicated
 *          pattern (i mod offset; offset 0 writes zeros, :309-314).
 *  rules   Everything near the end of the block, length chains, segment
 *          crossings and malformed input goes through lz4e_g_slow(), which
 *          applies the reference's checks one by one.
 */
#ifndef LZ4E_DECOMPRESS_GROUP_CUH
#define LZ4E_DECOMPRESS_GROUP_CUH

#include "lz4e_decompress_body.cuh"

#define LZ4E_G_MAXSEG 32u /* destination segments per block kept in the map */

/* destination segment map of one block (shared memory, one per group) */
struct lz4e_omap {
	uint32_t start[LZ4E_G_MAXSEG + 2]; /* linear position of segment k's byte 0 */
	uint64_t bias[LZ4E_G_MAXSEG];	   /* address of linear byte p = bias[k] + p */
};

/* every long copy of this decoder goes through ONE out-of-line instance of
 * lz4e_copy, so the hot loop stays small enough for the instruction cache */
template <int GS>
LZ4E_DEV_NOINLINE void lz4e_g_copy_far(uint8_t *dst, const uint8_t *src, uint32_t len,
				       uint32_t gl)
{
	lz4e_copy<GS>(dst, src, len, gl);
}

/* i mod d for small operands (i < 2^16, d >= 1) without an integer division */
LZ4E_DEV uint32_t lz4e_mod_small(uint32_t i, uint32_t d, float rcp)
{
	const uint32_t q = (uint32_t)(((float)i + 0.5f) * rcp);

	return i - q * d;
}

template <int GS, bool STAGE>
struct lz4e_gstate {
	/* source: byte p of the compressed stream is inb[p] while p < in_hi */
	const uint8_t *inb;
	const struct lz4e_b200_dev_bvec *sv;
	uint32_t in_hi, iend;
	/* destination: linear byte p is ob[p] for o_lo <= p < o_hi, pb[p] for
	 * p_lo <= p < o_lo (the segment before) */
	uint8_t *ob, *pb;
	uint32_t o_lo, o_hi, p_lo, oend;
	uint32_t oseg, nseg;
	uint32_t ip, op, b;
	bool active, done, scatter;
	bool irregular; /* the next sequence needs lz4e_g_slow() */
};

/* the group's lanes count the segments that start at or before p */
template <int GS>
LZ4E_DEV uint32_t lz4e_omap_find(const lz4e_omap *m, uint32_t nseg, uint32_t p,
				 uint32_t gl, uint32_t gmask)
{
	uint32_t cnt = 0;

	for (uint32_t k0 = 0; k0 < nseg; k0 += GS) {
		const uint32_t k = k0 + gl;
		const bool le = k < nseg && m->start[k] <= p;

		cnt += (uint32_t)lz4e_popc(lz4e_ballot(gmask, le));
	}
	return cnt - 1u; /* start[0] == 0 <= p */
}

/* make segment s the write cursor's segment */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_g_set_seg(lz4e_gstate<GS, STAGE> &g, const lz4e_omap *m, uint32_t s)
{
	g.oseg = s;
	g.o_lo = m->start[s];
	g.o_hi = lz4e_min(m->start[s + 1u], g.oend);
	g.ob = (uint8_t *)(uintptr_t)m->bias[s];
	LZ4E_ASSUME_GLOBAL(g.ob);
	if (s) {
		g.p_lo = m->start[s - 1u];
		g.pb = (uint8_t *)(uintptr_t)m->bias[s - 1u];
	} else {
		g.p_lo = 0;
		g.pb = g.ob;
	}
	LZ4E_ASSUME_GLOBAL(g.pb);
}

/* address of linear output byte p (p < op) and the end of its segment */
template <int GS, bool STAGE>
LZ4E_DEV uint8_t *lz4e_g_locate(const lz4e_gstate<GS, STAGE> &g, const lz4e_omap *m,
				uint32_t p, uint32_t &hi, uint32_t gl, uint32_t gmask)
{
	if (p >= g.o_lo) {
		hi = g.o_hi;
		return g.ob + p;
	}
	if (p >= g.p_lo) {
		hi = g.o_lo;
		return g.pb + p;
	}
	const uint32_t s = lz4e_omap_find<GS>(m, g.nseg, p, gl, gmask);
	uint8_t *q = (uint8_t *)(uintptr_t)m->bias[s] + p;

	LZ4E_ASSUME_GLOBAL(q);
	hi = m->start[s + 1u];
	return q;
}

/*
 * Match of len bytes at distance off inside ONE segment, source and
 * destination: o[i] = o[i - off] (all of [o - off, o + len) addressable from
 * o).  The group's earlier stores must be visible (barrier before the call).
 */
template <int GS>
LZ4E_DEV void lz4e_g_match_inseg(uint8_t *o, uint32_t off, uint32_t len, uint32_t gl,
				 uint32_t gmask)
{
	if (off == 0u) { /* lz4e_decompress.c:309-314 */
		for (uint32_t i = gl; i < len; i += GS)
			o[i] = 0;
	} else if (off >= len) {
		if (len >= 32u)
			lz4e_g_copy_far<GS>(o, o - off, len, gl);
		else
			for (uint32_t i = gl; i < len; i += GS)
				o[i] = (o - off)[i];
	} else if (len < 96u || off >= 48u) {
		if (off >= (uint32_t)GS) {
			/* every pass of GS bytes reads bytes written before it */
			for (uint32_t i0 = 0; i0 < len; i0 += GS) {
				if (i0)
					lz4e_syncwarp(gmask);
				if (i0 + gl < len)
					o[i0 + gl] = (o - off)[i0 + gl];
			}
		} else {
			const uint8_t *ms = o - off;
			uint32_t j = gl % off;
			const uint32_t dj = (uint32_t)GS % off;

			for (uint32_t i = gl; i < len; i += GS) {
				o[i] = ms[j];
				j += dj;
				if (j >= off)
					j -= off;
			}
		}
	} else {
		/* long pattern of period off: whole periods byte-wise until at
		 * least 32 bytes exist, then keep doubling with vector copies */
		const uint8_t *ms = o - off;
		uint32_t have = off * ((31u + off) / off);
		uint32_t j = gl % off;
		const uint32_t dj = (uint32_t)GS % off;

		for (uint32_t i = gl; i < have; i += GS) {
			o[i] = ms[j];
			j += dj;
			while (j >= off)
				j -= off;
		}
		while (have < len) {
			uint32_t chunk = len - have;

			if (chunk > have)
				chunk = have;
			lz4e_syncwarp(gmask);
			lz4e_g_copy_far<GS>(o + have, o, chunk, gl);
			have += chunk;
		}
	}
}

/* match of any shape: pieces are cut at destination and source segment ends */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_g_match(lz4e_gstate<GS, STAGE> &g, const lz4e_omap *m, uint32_t op,
			   uint32_t off, uint32_t len, uint32_t gl, uint32_t gmask)
{
	uint32_t done = 0;

	lz4e_syncwarp(gmask);
	if (STAGE || g.nseg <= 1u) {
		lz4e_g_match_inseg<GS>(g.ob + op, off, len, gl, gmask);
		return;
	}
	while (done < len) {
		const uint32_t dpos = op + done;

		if (dpos >= g.o_hi)
			lz4e_g_set_seg(g, m, g.oseg + 1u);
		uint32_t n = lz4e_min(len - done, g.o_hi - dpos);
		uint8_t *o = g.ob + dpos;

		if (off <= dpos - g.o_lo) {
			/* the source lies in this segment too */
			if (done)
				lz4e_syncwarp(gmask);
			lz4e_g_match_inseg<GS>(o, off, n, gl, gmask);
		} else {
			uint32_t shi;
			const uint8_t *s = lz4e_g_locate(g, m, dpos - off, shi, gl, gmask);

			n = lz4e_min(n, lz4e_min(shi - (dpos - off), off));
			if (done)
				lz4e_syncwarp(gmask);
			lz4e_g_copy_far<GS>(o, s, n, gl); /* n <= off: no overlap */
		}
		done += n;
	}
}

/* forward-only source view: make the segment holding position p current */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_g_in_seek(lz4e_gstate<GS, STAGE> &g, uint32_t p)
{
	while (p >= g.in_hi && g.in_hi < g.iend) {
		const uint32_t lo = g.in_hi;

		g.sv++;
		g.in_hi = lo + g.sv->len;
		if (g.in_hi > g.iend || g.in_hi < lo)
			g.in_hi = g.iend;
		g.inb = (const uint8_t *)(uintptr_t)(g.sv->addr - lo);
		LZ4E_ASSUME_GLOBAL(g.inb);
	}
}

template <int GS, bool STAGE>
LZ4E_DEV uint32_t lz4e_g_rd8(const lz4e_gstate<GS, STAGE> &g, uint32_t p)
{
	if (p < g.in_hi)
		return g.inb[p];
	/* the byte lies in a later segment */
	const struct lz4e_b200_dev_bvec *w = g.sv + 1;
	uint32_t base = g.in_hi;

	for (;;) {
		if (p - base < w->len) {
			const uint8_t *q = (const uint8_t *)(uintptr_t)(w->addr + (p - base));

			LZ4E_ASSUME_GLOBAL(q);
			return *q;
		}
		base += w->len;
		w++;
	}
}

/* literal run of len bytes from compressed position ip to linear output op;
 * pieces are cut at source and destination segment ends */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_g_literals(lz4e_gstate<GS, STAGE> &g, const lz4e_omap *m, uint32_t ip,
			      uint32_t op, uint32_t len, uint32_t gl)
{
	uint32_t done = 0;

	while (done < len) {
		const uint32_t dpos = op + done, spos = ip + done;

		if (!STAGE && dpos >= g.o_hi)
			lz4e_g_set_seg(g, m, g.oseg + 1u);
		lz4e_g_in_seek(g, spos);
		uint32_t n = lz4e_min(len - done, g.in_hi - spos);

		if (!STAGE)
			n = lz4e_min(n, g.o_hi - dpos);
		if (n < 16u) {
			if (gl < n)
				g.ob[dpos + gl] = g.inb[spos + gl];
			if (GS < 16 && gl + GS < n)
				g.ob[dpos + gl + GS] = g.inb[spos + gl + GS];
			if (GS < 8) {
				if (gl + 2 * GS < n)
					g.ob[dpos + gl + 2 * GS] = g.inb[spos + gl + 2 * GS];
				if (gl + 3 * GS < n)
					g.ob[dpos + gl + 3 * GS] = g.inb[spos + gl + 3 * GS];
			}
		} else {
			lz4e_g_copy_far<GS>(g.ob + dpos, g.inb + spos, n, gl);
		}
		done += n;
	}
}

/* result of a finished block; staged or scratch output goes home */
template <int GS, bool STAGE>
LZ4E_DEV void lz4e_g_finish(const lz4e_decompress_args &a, lz4e_gstate<GS, STAGE> &g,
			    int32_t ret, uint32_t gl, uint32_t gmask)
{
	lz4e_syncwarp(gmask);
	if ((STAGE || g.scatter) && ret > 0) {
		lz4e_cursor c = lz4e_cursor_init(a.dst_bvecs, a.dst_iters[g.b]);

		const uint8_t *src = g.ob;
		uint32_t left = (uint32_t)ret;

		while (left) { /* linear staging / scratch -> destination segments */
			const uint32_t vlen = c.v->len;
			const uint32_t take = lz4e_min(vlen - c.done, left);

			lz4e_g_copy_far<GS>(lz4e_cursor_addr(c), src, take, gl);
			src += take;
			left -= take;
			c.done += take;
			if (c.done == vlen) {
				c.v++;
				c.done = 0;
			}
		}
	}
	if (gl == 0)
		a.rets[g.b] = ret;
	g.active = false;
	lz4e_syncwarp(gmask); /* staging / map are reused by the next block */
}

/* take the next block of the batch: views, segment map, special cases */
template <int GS, bool STAGE>
LZ4E_DEV_NOINLINE void lz4e_g_open(const lz4e_decompress_args &a, lz4e_gstate<GS, STAGE> &g,
			  lz4e_omap *m, uint8_t *stage, uint32_t slot, uint32_t gl,
			  uint32_t gmask, int gleader)
{
	uint32_t b = 0;

	if (gl == 0)
		b = lz4e_atomic_add(a.work_counter, 1u);
	b = lz4e_shfl(gmask, b, gleader);
	if (b >= a.nblocks) {
		g.done = true;
		return;
	}
	if (a.order)
		b = a.order[b];
	const struct lz4e_b200_dev_iter si = a.src_iters[b];
	const struct lz4e_b200_dev_iter di = a.dst_iters[b];

	g.b = b;
	g.iend = si.bi_size;
	g.oend = di.bi_size;
	g.ip = g.op = 0;
	g.scatter = false;
	if (si.bi_size > (16u << 20) || di.bi_size > 0x7fffffffu ||
	    (STAGE && di.bi_size > a.stage_cap)) {
		if (gl == 0) {
			lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
			a.rets[b] = -1;
		}
		return;
	}
	g.sv = a.src_bvecs + si.first + si.bi_idx;
	if (si.bi_size) {
		g.in_hi = lz4e_min(g.sv->len - si.bi_bvec_done, si.bi_size);
		g.inb = (const uint8_t *)(uintptr_t)(g.sv->addr + si.bi_bvec_done);
	} else {
		g.in_hi = 0;
	}
	/* special cases, lz4e_decompress.c:113-120 */
	if (g.oend == 0 || g.iend == 0) {
		int32_t r = -1;

		if (g.oend == 0 && g.iend == 1 && g.inb[0] == 0)
			r = 0;
		if (gl == 0)
			a.rets[b] = r;
		return;
	}
	const struct lz4e_b200_dev_bvec *dv = a.dst_bvecs + di.first + di.bi_idx;
	const uint32_t first_len = dv->len - di.bi_bvec_done;

	g.nseg = 1;
	g.oseg = 0;
	g.o_lo = g.p_lo = 0;
	g.o_hi = g.oend;
	if (STAGE) {
		/* congruent mod 16 with the first destination segment, so the final
		 * scatter is vector stores */
		g.ob = stage + ((uint32_t)(dv->addr + di.bi_bvec_done) & 15u);
		g.pb = g.ob;
	} else if (first_len >= di.bi_size) {
		g.ob = (uint8_t *)(uintptr_t)(dv->addr + di.bi_bvec_done);
		LZ4E_ASSUME_GLOBAL(g.ob);
		g.pb = g.ob;
	} else {
		uint32_t pos = 0, k = 0;

		while (pos < di.bi_size && k < LZ4E_G_MAXSEG) {
			const uint32_t len = k ? dv[k].len : first_len;
			const uint64_t ad = k ? dv[k].addr : dv->addr + di.bi_bvec_done;

			if (gl == (k & (GS - 1))) {
				m->start[k] = pos;
				m->bias[k] = ad - pos;
			}
			pos += len;
			k++;
		}
		if (pos < di.bi_size) {
			/* more segments than the map holds: decode into a linear
			 * slot and scatter at the end */
			if (!a.scratch || di.bi_size > a.scratch_stride) {
				if (gl == 0) {
					lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
					a.rets[b] = -1;
				}
				return;
			}
			g.ob = a.scratch + (size_t)slot * a.scratch_stride;
			LZ4E_ASSUME_GLOBAL(g.ob);
			g.pb = g.ob;
			g.scatter = true;
		} else {
			if (gl == 0) {
				m->start[k] = pos;
				m->start[k + 1u] = pos;
			}
			g.nseg = k;
			lz4e_syncwarp(gmask);
			lz4e_g_set_seg(g, m, 0);
		}
	}
	g.active = true;
	g.irregular = false;
}

/*
 * One sequence by the book (lz4e_decompress.c:150-446): every end-of-block
 * rule, length chains, literals and matches that cross segments.
 */
template <int GS, bool STAGE>
LZ4E_DEV_NOINLINE void lz4e_g_slow(const lz4e_decompress_args &a, lz4e_gstate<GS, STAGE> &g,
			  lz4e_omap *m, uint32_t gl, uint32_t gmask)
{
	const int32_t iend = (int32_t)g.iend, oend = (int32_t)g.oend;
	uint32_t ip = g.ip, op = g.op;

	lz4e_g_in_seek(g, ip);
	const uint32_t token = lz4e_g_rd8(g, ip++);
	uint32_t len = token >> 4;
	const bool shortcut = len != 15u && (int32_t)ip < iend - 16 &&
			      (int32_t)op <= oend - 32; /* :150-156 */

	if (!shortcut) {
		if (len == 15u) { /* :194-220 */
			uint32_t s;

			if ((int32_t)ip >= iend - 15) {
				lz4e_g_finish(a, g, -(int32_t)ip - 1, gl, gmask);
				return;
			}
			do {
				s = lz4e_g_rd8(g, ip++);
				len += s;
			} while ((int32_t)ip < iend - 15 && s == 255u);
		}
		/* :222-266 */
		if ((int64_t)op + len > (int64_t)oend - 12 ||
		    (int64_t)ip + len > (int64_t)iend - 8) {
			if ((int64_t)ip + len != (int64_t)iend ||
			    (int64_t)op + len > (int64_t)oend) {
				lz4e_g_finish(a, g, -(int32_t)ip - 1, gl, gmask);
				return;
			}
			lz4e_g_literals(g, m, ip, op, len, gl);
			lz4e_g_finish(a, g, (int32_t)(op + len), gl, gmask);
			return;
		}
	}
	lz4e_g_literals(g, m, ip, op, len, gl);
	ip += len;
	op += len;

	/* offset, :280-283 */
	const uint32_t offset = lz4e_g_rd8(g, ip) | (lz4e_g_rd8(g, ip + 1) << 8);
	uint32_t mlen = token & 15u;

	ip += 2;
	if (offset > op) { /* :299-302 */
		lz4e_g_finish(a, g, -(int32_t)ip - 1, gl, gmask);
		return;
	}
	const bool full_shortcut = shortcut && mlen != 15u && offset >= 8u;

	if (mlen == 15u) { /* :316-334 */
		uint32_t s;

		do {
			s = lz4e_g_rd8(g, ip++);
			if ((int32_t)ip > iend - 5) {
				lz4e_g_finish(a, g, -(int32_t)ip - 1, gl, gmask);
				return;
			}
			mlen += s;
		} while (s == 255u);
	}
	mlen += 4u;
	if (!full_shortcut && (int64_t)op + mlen > (int64_t)oend - 5) { /* :420-431 */
		lz4e_g_finish(a, g, -(int32_t)ip - 1, gl, gmask);
		return;
	}
	lz4e_g_match(g, m, op, offset, mlen, gl, gmask);
	g.ip = ip;
	g.op = op + mlen;
}

/* the head of a match (its first 2 * GS bytes) waiting for its stores: lane gl
 * holds bytes gl and gl + GS */
struct lz4e_gpend {
	uint32_t pos; /* linear position of the match's first byte */
	uint32_t n;   /* bytes held, 0 = nothing pending */
	uint32_t v0, v1;
};

#define LZ4E_G_FLUSH(pd)                                             \
	do {                                                         \
		if (gl < (pd).n)                                     \
			ob[(pd).pos + gl] = (uint8_t)(pd).v0;        \
		if (gl + GS < (pd).n)                                \
			ob[(pd).pos + gl + GS] = (uint8_t)(pd).v1;   \
		(pd).n = 0;                                          \
	} while (0)

/*
 * The hot loop: REGULAR sequences only, all state in registers, no calls.
 * Regular = at most one extension byte per length (literals < 270, match < 274
 * bytes), at least 15 input bytes behind the offset field (rules out the input
 * checks at lz4e_decompress.c:197-200, :226-227, :322) and 12 output bytes left
 * after the match (:226, :425-431), so the shortcut predicate of :150-156 makes
 * no difference; literals inside the current source segment; destination inside
 * the current destination segment; match source inside ONE segment (the current
 * one, the one before, or any earlier one found through the map).
 * Every group of the warp advances by one sequence per step.  The loop ends
 * after the first step in which some group has no regular sequence (or no
 * block): that group is marked and the caller handles it out of line.
 */
template <int GS, bool STAGE, bool PIPE>
LZ4E_DEV void lz4e_g_hot(lz4e_gstate<GS, STAGE> &g, const lz4e_omap *m, uint32_t gl,
			 uint32_t gmask)
{
	const uint8_t *inb = g.inb;
	uint8_t *ob = g.ob, *pb = g.pb;
	const uint32_t o_lo = g.o_lo, p_lo = g.p_lo, nseg = g.nseg;
	const bool idle = g.done;
	const bool live = g.active && !g.done;
	/* last input position a regular sequence may start at; output limit */
	const uint32_t ihi = lz4e_min(g.in_hi, g.iend);
	const uint32_t lim = g.oend >= 12u ? lz4e_min(g.o_hi, g.oend - 12u) : 0u;
	uint32_t ip = g.ip, op = g.op;
	lz4e_gpend pend[2];

	/* (staged output is addressed generically: declaring the shared space on
	 * pointers that round-trip through the out-of-line state made nvcc 12.9
	 * prune the loop) */
	if (!STAGE) {
		LZ4E_ASSUME_GLOBAL(ob);
		LZ4E_ASSUME_GLOBAL(pb);
	}
	LZ4E_ASSUME_GLOBAL(inb);
	pend[0].n = pend[1].n = 0;
	pend[0].pos = pend[1].pos = pend[0].v0 = pend[0].v1 = pend[1].v0 = pend[1].v1 = 0;
	if (lz4e_ballot(LZ4E_FULL, live) == 0) /* nothing to decode (any more) */
		return;

	for (;;) {
#pragma unroll
	for (int ph = 0; ph < 2; ph++) {
		lz4e_gpend &old = pend[ph ^ 1]; /* match head of the previous step */
		lz4e_gpend &cur = pend[ph];	/* match head of this step */

		/* ---- parse: straight-line code, the few branches are warp-uniform ---- */
		const bool can = live && ip + 20u <= ihi;
		uint32_t lo = 0, hi = 0;

		if (can) {
			const uint8_t *p = inb + ip;
			const uint8_t *al = (const uint8_t *)((uintptr_t)p & ~(uintptr_t)3);
			const uint32_t sh = ((uint32_t)(uintptr_t)p & 3u) * 8u;
			const uint32_t w0 = lz4e_ld32a(al), w1 = lz4e_ld32a(al + 4),
				       w2 = lz4e_ld32a(al + 8);

			LZ4E_ASSUME_GLOBAL(al);
			lo = lz4e_funnel_r(w0, w1, sh);
			hi = lz4e_funnel_r(w1, w2, sh);
		}
		const uint32_t l0 = (lo >> 4) & 15u;
		const uint32_t ml4 = lo & 15u;
		const uint32_t e1 = (lo >> 8) & 0xffu;
		const bool lext = l0 == 15u;
		const uint32_t lit = l0 + (lext ? e1 : 0u);
		const uint32_t lp = ip + 1u + (lext ? 1u : 0u);
		const uint32_t fp = lp + lit; /* offset field */
		bool ok = can && !(lext && e1 == 255u) && fp + 15u <= ihi;
		/* offset + first match-length byte: inside the window for up to 4
		 * literals (a 64-bit shift: the funnel shift takes its count mod 32) */
		uint32_t ow = (uint32_t)((((uint64_t)hi << 32) | lo) >> ((8u + 8u * lit) & 63u));
		const bool second = ok && lit > 4u;

		if (lz4e_ballot(LZ4E_FULL, second)) {
			if (second)
				ow = lz4e_ld32u(inb + fp);
		}
		const uint32_t off = ow & 0xffffu;
		const uint32_t e2 = (ow >> 16) & 0xffu;
		const bool mext = ml4 == 15u;
		const uint32_t mlen = ml4 + 4u + (mext ? e2 : 0u);
		const uint32_t mo = op + lit; /* match destination */
		const uint32_t sp = mo - off; /* match source */
		/* (an impossible offset must not form a wild pointer: the address
		 * space of ms is declared to the compiler below) */
		const uint8_t *ms = ob + (off <= mo ? sp : 0u);

		ok = ok && !(mext && e2 == 255u) && off != 0u && off <= mo && mo + mlen <= lim;
		if (!STAGE) {
			/* source before the current segment: all of it must lie in ONE
			 * earlier segment (then it cannot overlap the destination) */
			const bool back = ok && sp < o_lo;

			if (lz4e_ballot(LZ4E_FULL, back)) {
				if (back) {
					uint32_t shi = o_lo;

					if (sp >= p_lo) {
						ms = pb + sp;
					} else {
						const uint32_t sg =
							lz4e_omap_find<GS>(m, nseg, sp, gl, gmask);

						ms = (const uint8_t *)(uintptr_t)m->bias[sg] + sp;
						shi = m->start[sg + 1u];
					}
					ok = off >= mlen && sp + mlen <= shi;
				}
			}
		}
		if (!STAGE)
			LZ4E_ASSUME_GLOBAL(ms);
		const uint32_t f_lit = ok ? lit : 0u, f_mlen = ok ? mlen : 0u;
		const uint32_t f_off = ok ? off : 1u;
		const uint32_t f_win = lz4e_funnel_r(lo, hi, 8u);
		uint8_t *ld = ob + op;	       /* literal destination */
		uint8_t *md = ob + mo;	       /* match destination */
		const uint8_t *ls = inb + lp; /* literal source */
		const bool ovl = f_off < f_mlen;
		/* a group that is out of blocks just idles along */
		const bool stop = lz4e_ballot(LZ4E_FULL, !ok && !idle) != 0;

		if (ok) {
			ip = fp + 2u + (mext ? 1u : 0u);
			op = mo + mlen;
		}

		/* ---- copy phase: all 32 lanes, every group its own sequence
		 * (groups without one have f_lit == f_mlen == 0) ---- */
		if (PIPE) {
			/* stores of the previous match head that this match reads
			 * must go first (they become visible at the barrier below) */
			const bool dep = old.n && f_mlen && sp + (ovl ? f_off : f_mlen) > old.pos;

			if (lz4e_ballot(LZ4E_FULL, dep))
				LZ4E_G_FLUSH(old);
		}
		/* literals: up to 4 out of the parse window, longer runs from the
		 * compressed stream */
		if (gl < f_lit && f_lit <= 4u)
			ld[gl] = (uint8_t)(f_win >> (8u * gl));
		if (lz4e_ballot(LZ4E_FULL, f_lit > 4u)) {
			if (f_lit > 4u) {
				uint32_t i = gl;

				for (; i + GS < f_lit; i += 2 * GS) {
					const uint8_t x = ls[i], y = ls[i + GS];

					ld[i] = x;
					ld[i + GS] = y;
				}
				if (i < f_lit)
					ld[i] = ls[i];
			}
		}
		lz4e_syncwarp(LZ4E_FULL);
		{
			uint32_t j0 = gl, j1 = gl + GS;
			const uint32_t head = PIPE ? lz4e_min(f_mlen, 2u * GS) : 0u;
			float rcp = 1.0f;

			/* an overlapping match is a pattern of period off: byte i
			 * comes from source byte i mod off, which was final before
			 * the match began, so no pass depends on another */
			if (lz4e_ballot(LZ4E_FULL, ovl)) {
				if (ovl) {
					rcp = lz4e_rcpf((float)f_off);
					j0 = lz4e_mod_small(gl, f_off, rcp);
					j1 = lz4e_mod_small(gl + GS, f_off, rcp);
				}
			}
			if (PIPE) {
				cur.v0 = cur.v1 = 0;
				if (gl < head)
					cur.v0 = ms[j0];
				if (gl + GS < head)
					cur.v1 = ms[j1];
				cur.pos = mo;
				cur.n = head;
			}
			/* the rest of a long match (all of it without PIPE) */
			if (lz4e_ballot(LZ4E_FULL, f_mlen > head)) {
				if (f_mlen > head && !ovl) {
					uint32_t i = head + gl;

					for (; i + GS < f_mlen; i += 2 * GS) {
						const uint8_t x = ms[i], y = ms[i + GS];

						md[i] = x;
						md[i + GS] = y;
					}
					if (i < f_mlen)
						md[i] = ms[i];
				} else if (f_mlen > head) {
					const uint32_t dj = lz4e_mod_small(GS, f_off, rcp);
					uint32_t j = PIPE ? (j1 + dj >= f_off ? j1 + dj - f_off : j1 + dj)
							  : j0;

					for (uint32_t i = head + gl; i < f_mlen; i += GS) {
						md[i] = ms[j];
						j += dj;
						if (j >= f_off)
							j -= f_off;
					}
				}
			}
			if (PIPE) /* the previous head: its loads were issued a step ago */
				LZ4E_G_FLUSH(old);
		}
		if (stop) {
			if (PIPE)
				LZ4E_G_FLUSH(cur);
			g.ip = ip;
			g.op = op;
			g.irregular = live && !ok;
			return;
		}
	}
	}
}

/*
 * sm: this warp's shared memory: 32/GS staging areas of stage_cap + LZ4E_D_PAD
 * bytes (STAGE) or 32/GS segment maps; slot0: this warp's first scratch slot.
 */
template <int GS, bool STAGE, bool PIPE>
LZ4E_DEV void lz4e_decompress_groups(const lz4e_decompress_args &a, uint8_t *sm,
				     uint32_t slot0, uint32_t lane)
{
	const uint32_t gl = lane & (GS - 1);
	const uint32_t gi = lane / GS;
	const uint32_t gmask = (GS == 32 ? 0xffffffffu : ((1u << GS) - 1u)) << (gi * GS);
	const int gleader = (int)(gi * GS);
	lz4e_omap *m = STAGE ? nullptr : reinterpret_cast<lz4e_omap *>(sm) + gi;
	uint8_t *stage = STAGE ? sm + gi * (a.stage_cap + LZ4E_D_PAD) : nullptr;
	lz4e_gstate<GS, STAGE> g;

	/* valid pointers of the right address space even before the first block
	 * (the hot loop declares the spaces to the compiler) */
	g.inb = (const uint8_t *)a.src_bvecs;
	g.sv = a.src_bvecs;
	g.ob = g.pb = STAGE ? stage : (uint8_t *)a.rets;
	g.in_hi = g.iend = g.o_lo = g.o_hi = g.p_lo = g.oend = 0;
	g.oseg = 0;
	g.nseg = 1;
	g.ip = g.op = g.b = 0;
	g.active = g.done = g.scatter = g.irregular = false;

	for (;;) {
		if (lz4e_ballot(LZ4E_FULL, !g.done) == 0)
			return;
		/* ---- out of line: a new block, or one sequence by the book ---- */
		if (!g.done) {
			if (!g.active)
				lz4e_g_open(a, g, m, stage, slot0 + gi, gl, gmask, gleader);
			else if (g.irregular)
				lz4e_g_slow(a, g, m, gl, gmask);
		}
		lz4e_syncwarp(LZ4E_FULL);
		/* ---- regular sequences until some group needs the above again ---- */
		lz4e_g_hot<GS, STAGE, PIPE>(g, m, gl, gmask);
	}
}

#endif /* LZ4E_DECOMPRESS_GROUP_CUH */
