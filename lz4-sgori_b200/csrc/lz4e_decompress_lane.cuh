/*
 * lz4e_decompress_lane.cuh — the LANE-PER-BLOCK decoder: every lane of a warp
 * parses its own LZ4 block, all 32 lanes copy together.
 *
 * Semantics are those of lz4e_decompress_body.cuh (LZ4E_decompress_safe,
 * lz4e/lz4e_decompress.c:62-470 of the reference: same accepted streams, same
 * output, same -(input offset) - 1 for malformed ones, :458-459).  What differs
 * is how the work is laid out on the warp:
 *
 *   parse   The token chain of a block is serial (token -> lengths -> next
 *           token), so a group of lanes that parses ONE sequence together
 *           repeats the same ~50 instructions on every lane.  Here each lane
 *           parses the next sequence of ITS OWN block: one instruction stream
 *           yields 32 sequences.  The result is a descriptor
 *           {literal source, output position, literal length, offset, match
 *           length}; the parse itself moves no data.
 *   copy    The 32 descriptors belong to 32 different blocks, so they do not
 *           depend on each other.  Their bytes are cut into units of 8; a
 *           prefix sum over the unit counts gives every unit a slot, and in
 *           each round 4 units are moved by 8 lanes each (one byte per lane,
 *           8 contiguous bytes per unit: one 32-byte sector touched per unit).
 *           All literals first, then all matches, because a match may read
 *           the literals of its own sequence.  Overlapping matches use the
 *           replicated-pattern index (i mod offset), offset 0 writes zeros
 *           (lz4e_decompress.c:309-314).
 *   long    A literal run or match of LZ4E_L_WIDE bytes or more (and the
 *           gather/scatter of a block whose source or destination is not one
 *           contiguous segment) is posted and served by all 32 lanes with
 *           16-byte vectors at the top of the next iteration.
 *
 * Source and destination: a block whose compressed bytes lie in one segment
 * is read in place, otherwise the segments are first gathered into a linear
 * slot (args.in_scratch).  A destination that is one segment is written in
 * place, otherwise the block is decoded into a linear slot (args.scratch) and
 * scattered when it is complete.
 */
#ifndef LZ4E_DECOMPRESS_LANE_CUH
#define LZ4E_DECOMPRESS_LANE_CUH

#include "lz4e_decompress_body.cuh"

#define LZ4E_L_UNIT 8u
#define LZ4E_L_WIDE 512u /* runs of at least this many bytes go to the whole warp */
#define LZ4E_L_TAB 256u	 /* unit entries per warp; longer lists take several passes */
#ifndef LZ4E_L_BATCH
#define LZ4E_L_BATCH 4u /* rounds whose loads are in flight together */
#endif
#define LZ4E_L_PAD (LZ4E_L_BATCH * (32u / LZ4E_L_UNIT))
#define LZ4E_P_GATHER 8u
#define LZ4E_P_SCATTER 16u
#define LZ4E_L_SMALL 0x80000000u

/*
 * Per-warp shared memory of the copy phase: one 16-byte entry per unit.
 *   x, y  address of the unit's first output byte (y: bits 0-23 = address
 *         bits 32-55, bits 24-31 = bytes in the unit, 0..8; 0 = padding)
 *   literals: z, w = address of the unit's first source byte
 *   matches:  byte j of the unit is copied from D bytes back, with
 *             D = z + (j >= ((w >> 16) & 255) ? (w & 0xffff) : 0):
 *             z is the offset times the number of periods the unit's first
 *             byte lies past the match start (the replicated pattern of an
 *             overlapping match), w & 0xffff the offset, bits 16-23 the first
 *             byte of the unit that lies one period further.  Offsets below 8
 *             that overlap (several periods inside one unit, and offset 0 =
 *             zeros) are marked LZ4E_L_SMALL: z = index of the unit's first
 *             byte inside the match, the reader takes the remainder itself.
 */
struct lz4e_lane_smem {
	lz4e_u128 unit[LZ4E_L_TAB + LZ4E_L_PAD];
};

/* the 8 bytes at p (any alignment, little endian) from two aligned 8-byte
 * loads; reads up to 15 bytes past p */
LZ4E_DEV uint64_t lz4e_ld64u(const uint8_t *p)
{
	const uintptr_t a = (uintptr_t)p;
	const uint64_t *q = (const uint64_t *)(a & ~(uintptr_t)7);
	const uint32_t sh = (uint32_t)(a & 7u) * 8u;
	const uint64_t lo = q[0], hi = q[1];

	return sh ? (lo >> sh) | (hi << (64u - sh)) : lo;
}

/* inclusive prefix sum over the warp; *total = sum over all lanes */
LZ4E_DEV uint32_t lz4e_lane_scan(uint32_t v, uint32_t lane, uint32_t *total)
{
	for (uint32_t d = 1; d < 32u; d <<= 1) {
		const uint32_t t = lz4e_shfl(LZ4E_FULL, v, (int)(lane - d));

		if (lane >= d)
			v += t;
	}
	*total = lz4e_shfl(LZ4E_FULL, v, 31);
	return v;
}

/*
 * Move one run per lane (len < LZ4E_L_WIDE bytes, 0 = nothing) with all 32
 * lanes; the lane's units are first .. first + ceil(len / 8) - 1 of `total`
 * (total > 0).  MATCH: the run is dst[i] = dst[i - off] (pattern of period
 * off when off < len, zeros when off == 0); otherwise dst[i] = src[i].
 * Ends with a warp barrier.
 */
template <bool MATCH>
LZ4E_DEV void lz4e_lane_spread(lz4e_lane_smem *s, uint8_t *dst, const uint8_t *src,
			       uint32_t off, uint32_t len, uint32_t first,
			       uint32_t total, uint32_t lane)
{
	const uint32_t sub = lane & (LZ4E_L_UNIT - 1u);
	const uint32_t last = first + (len + LZ4E_L_UNIT - 1u) / LZ4E_L_UNIT;
	const uint64_t d0 = (uint64_t)(uintptr_t)dst;
	const uint64_t s0 = (uint64_t)(uintptr_t)src;

	for (uint32_t base = 0; base < total; base += LZ4E_L_TAB) {
		const uint32_t cnt = lz4e_min(total - base, LZ4E_L_TAB);
		const uint32_t lo = first > base ? first : base;
		const uint32_t hi = lz4e_min(last, base + LZ4E_L_TAB);

		/* ---- every lane lists the units of its run ---- */
		if (lo < hi) {
			uint32_t k = (lo - first) * LZ4E_L_UNIT;
			const bool small = MATCH && off < LZ4E_L_UNIT;
			uint32_t r = 0, dist = off;

			if (MATCH && !small && k) { /* later pass of a long list */
				r = k % off;
				dist = off * (k / off + 1u);
			}
			for (uint32_t u = lo; u < hi; u++, k += LZ4E_L_UNIT) {
				const uint32_t n = lz4e_min(len - k, LZ4E_L_UNIT);
				const uint64_t d = d0 + k;
				lz4e_u128 e;

				e.x = (uint32_t)d;
				e.y = ((uint32_t)(d >> 32) & 0xffffffu) | (n << 24);
				if (!MATCH) {
					e.z = (uint32_t)(s0 + k);
					e.w = (uint32_t)((s0 + k) >> 32);
				} else if (small && k + n > off) {
					e.z = k;
					e.w = off | LZ4E_L_SMALL;
				} else {
					const uint32_t cross = lz4e_min(off - r, LZ4E_L_UNIT);

					e.z = dist;
					e.w = off | (cross << 16);
					r += LZ4E_L_UNIT;
					if (r >= off) {
						r -= off;
						dist += off;
					}
				}
				lz4e_st128((uint8_t *)&s->unit[u - base], e);
			}
		}
		if (lane < LZ4E_L_PAD) { /* empty entries up to a whole batch */
			lz4e_u128 e = { 0, 0, 0, 0 };

			lz4e_st128((uint8_t *)&s->unit[cnt + lane], e);
		}
		lz4e_syncwarp(LZ4E_FULL);
		/*
		 * No unit reads what another unit of the same phase writes
		 * (different blocks; an overlapping match reads only bytes before
		 * its own start), so the loads of LZ4E_L_BATCH rounds are issued
		 * before their stores: the global-memory latency is paid once
		 * per batch, not once per round.
		 */
		for (uint32_t u0 = lane / LZ4E_L_UNIT; u0 < cnt; u0 += LZ4E_L_PAD) {
			uint8_t *wp[LZ4E_L_BATCH];
			uint32_t wv[LZ4E_L_BATCH];
			bool on[LZ4E_L_BATCH];

#pragma unroll
			for (uint32_t q = 0; q < LZ4E_L_BATCH; q++) {
				const lz4e_u128 e = lz4e_ld128((const uint8_t *)&s->unit[
					u0 + q * (32u / LZ4E_L_UNIT)]);
				uint8_t *o = (uint8_t *)(uintptr_t)(
					(((uint64_t)(e.y & 0xffffffu) << 32) | e.x) + sub);

				on[q] = sub < (e.y >> 24);
				wp[q] = o;
				wv[q] = 0;
				if (MATCH) {
					const uint8_t *p =
						o - (e.z + (sub >= ((e.w >> 16) & 255u) ? (e.w & 0xffffu)
										    : 0u));
					bool ld = on[q];

					if (e.w & LZ4E_L_SMALL) {
						const uint32_t f = e.w & 0xffffu;
						const uint32_t i = e.z + sub;

						if (f)
							p = o - i - f + i % f;
						else
							ld = false;
					}
					if (ld) {
						LZ4E_ASSUME_GLOBAL(p);
						wv[q] = *p;
					}
				} else if (on[q]) {
					const uint8_t *p = (const uint8_t *)(uintptr_t)(
						(((uint64_t)e.w << 32) | e.z) + sub);

					LZ4E_ASSUME_GLOBAL(p);
					wv[q] = *p;
				}
			}
#pragma unroll
			for (uint32_t q = 0; q < LZ4E_L_BATCH; q++)
				if (on[q]) {
					LZ4E_ASSUME_GLOBAL(wp[q]);
					*wp[q] = (uint8_t)wv[q];
				}
		}
		lz4e_syncwarp(LZ4E_FULL);
	}
}

struct lz4e_dlane {
	const uint8_t *in; /* linear compressed bytes */
	uint8_t *out;	   /* linear output */
	uint32_t b;
	int32_t iend, oend;
	uint32_t ip, op;
	bool active, done, scatter;
	uint32_t pend;
	const uint8_t *p_lsrc;
	uint32_t p_lop, p_llen, p_mop, p_mdist, p_mlen;
	int32_t p_ret;
};

/* take block b: resolve its source and destination, post a gather if the
 * compressed bytes are not contiguous.  slot = this lane's scratch slot. */
LZ4E_DEV void lz4e_dlane_open(const lz4e_decompress_args &a, lz4e_dlane &g,
			      uint32_t b, uint32_t slot)
{
	const struct lz4e_b200_dev_iter si = a.src_iters[b];
	const struct lz4e_b200_dev_iter di = a.dst_iters[b];

	g.b = b;
	g.iend = (int32_t)si.bi_size;
	g.oend = (int32_t)di.bi_size;
	g.ip = g.op = 0;
	g.scatter = false;
	if (si.bi_size > (16u << 20) || di.bi_size > 0x7fffffffu) {
		lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
		a.rets[b] = -1;
		return;
	}
	const struct lz4e_b200_dev_bvec *sv = a.src_bvecs + si.first + si.bi_idx;

	/* special cases, lz4e_decompress.c:113-120 */
	if (g.oend == 0 || g.iend == 0) {
		int32_t r = -1;

		if (g.oend == 0 && g.iend == 1) {
			const uint8_t *p =
				(const uint8_t *)(uintptr_t)(sv->addr + si.bi_bvec_done);

			LZ4E_ASSUME_GLOBAL(p);
			if (*p == 0)
				r = 0;
		}
		a.rets[b] = r;
		return;
	}
	const struct lz4e_b200_dev_bvec *dv = a.dst_bvecs + di.first + di.bi_idx;

	if (dv->len - di.bi_bvec_done >= di.bi_size) {
		g.out = (uint8_t *)(uintptr_t)(dv->addr + di.bi_bvec_done);
	} else {
		if (!a.scratch || di.bi_size > a.scratch_stride) {
			lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
			a.rets[b] = -1;
			return;
		}
		g.out = a.scratch + (size_t)slot * a.scratch_stride;
		g.scatter = true;
	}
	LZ4E_ASSUME_GLOBAL(g.out);
	if (sv->len - si.bi_bvec_done >= si.bi_size) {
		g.in = (const uint8_t *)(uintptr_t)(sv->addr + si.bi_bvec_done);
	} else {
		if (!a.in_scratch || si.bi_size > a.in_stride) {
			lz4e_atomic_or(a.faults, LZ4E_FAULT_RANGE);
			a.rets[b] = -1;
			return;
		}
		g.in = a.in_scratch + (size_t)slot * a.in_stride;
		g.pend = LZ4E_P_GATHER;
	}
	LZ4E_ASSUME_GLOBAL(g.in);
	g.active = true;
}

/* posted long copies, gathers and scatters, served by all 32 lanes */
LZ4E_DEV void lz4e_dlane_service(const lz4e_decompress_args &a, lz4e_dlane &g,
				 uint32_t lane)
{
	uint32_t pm = lz4e_ballot(LZ4E_FULL, g.pend != 0u);

	if (pm == 0)
		return;
	lz4e_syncwarp(LZ4E_FULL);
	do {
		const int L = lz4e_ffs(pm) - 1;
		const uint32_t what = lz4e_shfl(LZ4E_FULL, g.pend, L);
		uint8_t *out = (uint8_t *)(uintptr_t)lz4e_shfl64(
			(uint64_t)(uintptr_t)g.out, L);

		LZ4E_ASSUME_GLOBAL(out);
		if (what & LZ4E_P_GATHER) {
			uint8_t *in = (uint8_t *)(uintptr_t)lz4e_shfl64(
				(uint64_t)(uintptr_t)g.in, L);
			const uint32_t bb = lz4e_shfl(LZ4E_FULL, g.b, L);
			const uint32_t n = lz4e_shfl(LZ4E_FULL, (uint32_t)g.iend, L);
			lz4e_cursor c = lz4e_cursor_init(a.src_bvecs, a.src_iters[bb]);

			LZ4E_ASSUME_GLOBAL(in);
			lz4e_gather<32>(in, c, n, lane);
		}
		if (what & LZ4E_P_LIT) {
			const uint8_t *lsrc = (const uint8_t *)(uintptr_t)lz4e_shfl64(
				(uint64_t)(uintptr_t)g.p_lsrc, L);
			const uint32_t lop = lz4e_shfl(LZ4E_FULL, g.p_lop, L);
			const uint32_t llen = lz4e_shfl(LZ4E_FULL, g.p_llen, L);

			LZ4E_ASSUME_GLOBAL(lsrc);
			lz4e_copy<32>(out + lop, lsrc, llen, lane);
		}
		if (what & LZ4E_P_MATCH) {
			const uint32_t mop = lz4e_shfl(LZ4E_FULL, g.p_mop, L);
			const uint32_t mdist = lz4e_shfl(LZ4E_FULL, g.p_mdist, L);
			const uint32_t mlen = lz4e_shfl(LZ4E_FULL, g.p_mlen, L);

			lz4e_syncwarp(LZ4E_FULL);
			lz4e_wide_match(out + mop, mdist, mlen, lane);
		}
		if (what & LZ4E_P_SCATTER) {
			const uint32_t bb = lz4e_shfl(LZ4E_FULL, g.b, L);
			const uint32_t n = lz4e_shfl(LZ4E_FULL, (uint32_t)g.p_ret, L);
			lz4e_cursor c = lz4e_cursor_init(a.dst_bvecs, a.dst_iters[bb]);

			lz4e_syncwarp(LZ4E_FULL);
			lz4e_scatter<32>(c, out, n, lane);
		}
		pm &= pm - 1u;
	} while (pm);
	lz4e_syncwarp(LZ4E_FULL);
	if (g.pend) {
		if (g.pend & LZ4E_P_FINISH) {
			a.rets[g.b] = g.p_ret;
			g.active = false;
		}
		g.pend = 0;
	}
}

/*
 * sm: this warp's lz4e_lane_smem; slot0: index of this warp's first scratch
 * slot (one slot per lane).
 */
LZ4E_DEV void lz4e_decompress_lanes(const lz4e_decompress_args &a,
				    lz4e_lane_smem *sm, uint32_t slot0,
				    uint32_t lane)
{
	lz4e_dlane g;

	g.in = nullptr;
	g.out = nullptr;
	g.b = 0;
	g.iend = g.oend = 0;
	g.ip = g.op = 0;
	g.active = g.done = g.scatter = false;
	g.pend = 0;
	g.p_lsrc = nullptr;
	g.p_lop = g.p_llen = g.p_mop = g.p_mdist = g.p_mlen = 0;
	g.p_ret = 0;
	for (;;) {
		if (lz4e_ballot(LZ4E_FULL, !g.done) == 0)
			break;
		lz4e_dlane_service(a, g, lane);
		/* ---- hand out blocks to the lanes that have none ---- */
		{
			const bool want = !g.active && !g.done;
			const uint32_t need = lz4e_ballot(LZ4E_FULL, want);

			if (need) {
				const int leader = lz4e_ffs(need) - 1;
				uint32_t base = 0;

				if ((int)lane == leader)
					base = lz4e_atomic_add(a.work_counter,
							       (uint32_t)lz4e_popc(need));
				base = lz4e_shfl(LZ4E_FULL, base, leader);
				if (want) {
					const uint32_t b =
						base + (uint32_t)lz4e_popc(need & ((1u << lane) - 1u));

					if (b >= a.nblocks)
						g.done = true;
					else
						lz4e_dlane_open(a, g, a.order ? a.order[b] : b,
								slot0 + lane);
				}
			}
		}
		/* ---- parse one sequence of this lane's block ---- */
		uint32_t c_lit = 0, c_mlen = 0, c_off = 1;
		const uint8_t *c_lsrc = nullptr;
		uint8_t *c_o = nullptr;
		bool fin = false;
		int32_t ret = 0;

		if (g.active && g.pend == 0u)
		do {
			const int32_t iend = g.iend, oend = g.oend;
			const uint8_t *inb = g.in;
			uint32_t ip = g.ip, op = g.op;

			/*
			 * Common path (same predicate as lz4e_decompress_body.cuh): at
			 * most one extension byte per length, far enough from both ends
			 * that none of the reference's end-of-block rules can fire.
			 */
			if ((int32_t)ip + 16 <= iend) {
				/*
				 * The 8 bytes at ip in one register pair (two aligned 8-byte
				 * loads): token, literal-length byte and — when the literal
				 * run is at most 3-4 bytes, which is nearly always — offset
				 * and match-length byte as well.  Every lane reads its own
				 * stream, so a byte load is 32 sectors for 32 bytes: five of
				 * them per sequence were half of this kernel's LSU traffic.
				 */
				const uint64_t w = lz4e_ld64u(inb + ip);
				const uint32_t tok = (uint32_t)w & 0xffu;
				const uint32_t e1 = ((uint32_t)w >> 8) & 0xffu;
				const uint32_t lext = (tok >> 4) == 15u;
				const uint32_t lit = (tok >> 4) + (lext ? e1 : 0u);
				const uint32_t lp = ip + 1u + lext;
				const uint32_t fp = lp + lit;

				if (!(lext && e1 == 255u) && (int32_t)fp + 16 <= iend) {
					const uint32_t rel = fp - ip;
					const uint64_t w2 = rel <= 5u ? w >> (8u * rel)
								      : lz4e_ld64u(inb + fp);
					const uint32_t off = (uint32_t)w2 & 0xffffu;
					const uint32_t e2 = ((uint32_t)w2 >> 16) & 0xffu;
					const uint32_t mext = (tok & 15u) == 15u;
					const uint32_t mlen = (tok & 15u) + 4u + (mext ? e2 : 0u);

					if (!(mext && e2 == 255u) && off != 0u && off <= op + lit &&
					    lit < LZ4E_L_WIDE && mlen < LZ4E_L_WIDE &&
					    (int64_t)op + lit + mlen <= (int64_t)oend - 12) {
						c_o = g.out + op;
						c_lsrc = inb + lp;
						c_lit = lit;
						c_off = off;
						c_mlen = mlen;
						g.ip = fp + 2u + mext;
						g.op = op + lit + mlen;
						break;
					}
				}
			}
			/* ---- general path: the reference's rules one by one ---- */
			const uint32_t token = inb[ip++];
			uint32_t len = token >> 4;
			const bool shortcut = len != 15u && (int32_t)ip < iend - 16 &&
					      (int32_t)op <= oend - 32; /* :150-156 */

			fin = true; /* until the sequence turns out to be complete */
			if (!shortcut) {
				if (len == 15u) { /* :194-220 */
					uint32_t s;

					if ((int32_t)ip >= iend - 15) {
						ret = -(int32_t)ip - 1;
						break;
					}
					do {
						s = inb[ip++];
						len += s;
					} while ((int32_t)ip < iend - 15 && s == 255u);
				}
				/* :222-266 */
				if ((int64_t)op + len > (int64_t)oend - 12 ||
				    (int64_t)ip + len > (int64_t)iend - 8) {
					if ((int64_t)ip + len != (int64_t)iend ||
					    (int64_t)op + len > (int64_t)oend) {
						ret = -(int32_t)ip - 1;
						break;
					}
					/* last literals: the block ends here */
					c_o = g.out + op;
					c_lsrc = inb + ip;
					c_lit = len;
					ret = (int32_t)(op + len);
					break;
				}
			}
			c_o = g.out + op;
			c_lsrc = inb + ip;
			c_lit = len;
			ip += len;
			op += len;
			/* offset, :280-283 */
			const uint32_t offset = (uint32_t)inb[ip] | ((uint32_t)inb[ip + 1u] << 8);
			uint32_t mlen = token & 15u;

			ip += 2;
			if (offset > op) { /* :299-302 */
				ret = -(int32_t)ip - 1;
				break;
			}
			const bool full_shortcut = shortcut && mlen != 15u && offset >= 8u;

			if (mlen == 15u) { /* :316-334 */
				uint32_t s;
				bool bad = false;

				do {
					s = inb[ip++];
					if ((int32_t)ip > iend - 5) {
						bad = true;
						break;
					}
					mlen += s;
				} while (s == 255u);
				if (bad) {
					ret = -(int32_t)ip - 1;
					break;
				}
			}
			mlen += 4u;
			if (!full_shortcut &&
			    (int64_t)op + mlen > (int64_t)oend - 5) { /* :420-431 */
				ret = -(int32_t)ip - 1;
				break;
			}
			fin = false;
			c_off = offset;
			c_mlen = mlen;
			g.ip = ip;
			g.op = op + mlen;
		} while (0);

		/* runs too long for the unit table go to the whole warp */
		bool wide = false;

		if (c_lit >= LZ4E_L_WIDE || c_mlen >= LZ4E_L_WIDE) {
			wide = true;
			g.pend = (c_lit ? LZ4E_P_LIT : 0u) | (c_mlen ? LZ4E_P_MATCH : 0u);
			g.p_lsrc = c_lsrc;
			g.p_lop = (uint32_t)(c_o - g.out);
			g.p_llen = c_lit;
			g.p_mop = g.p_lop + c_lit;
			g.p_mdist = c_off;
			g.p_mlen = c_mlen;
			c_lit = c_mlen = 0;
		}
		/* ---- copy phase: all literals, then all matches ---- */
		{
			const uint32_t nul = (c_lit + LZ4E_L_UNIT - 1u) / LZ4E_L_UNIT;
			const uint32_t num = (c_mlen + LZ4E_L_UNIT - 1u) / LZ4E_L_UNIT;
			uint32_t tot;
			const uint32_t incl = lz4e_lane_scan(nul | (num << 16), lane, &tot);

			if (tot & 0xffffu)
				lz4e_lane_spread<false>(sm, c_o, c_lsrc, 0u, c_lit,
							(incl & 0xffffu) - nul, tot & 0xffffu, lane);
			if (tot >> 16)
				lz4e_lane_spread<true>(sm, c_o + c_lit, nullptr, c_off, c_mlen,
						       (incl >> 16) - num, tot >> 16, lane);
		}
		if (fin) {
			const bool sc = g.scatter && ret > 0;

			if (wide || sc) {
				g.pend |= (sc ? LZ4E_P_SCATTER : 0u) | LZ4E_P_FINISH;
				g.p_ret = ret;
			} else {
				a.rets[g.b] = ret;
				g.active = false;
			}
		}
	}
}

#endif /* LZ4E_DECOMPRESS_LANE_CUH */
