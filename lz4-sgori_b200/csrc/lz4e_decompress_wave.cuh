/*
 * lz4e_decompress_wave.cuh — the WAVE decoder: one warp decodes one LZ4 block,
 * 32 sequences at a time.
 *
 * Semantics: LZ4E_decompress_safe (lz4e/lz4e_decompress.c:462-470 of the
 * reference), i.e. LZ4E_decompress_generic (:62-460) with endOnInputSize,
 * decode_full_block, noDict, lowPrefix == dest — the same accepted streams, the
 * same output, the same -(input offset) - 1 for malformed ones (:458-459) —
 * over scatter-gather sources and destinations.  Everything the reference
 * decides near the end of a block, on length chains and on malformed input is
 * decided by lz4e_g_slow() of lz4e_decompress_group.cuh ("one sequence by the
 * book"); this file only adds a faster way through REGULAR sequences (the
 * definition is the one of lz4e_g_hot(): at most one extension byte per
 * length, 15 input bytes behind the offset field, 12 output bytes behind the
 * match, all inside the current source and destination segments).
 *
 * The serial token chain of a block (token -> lengths -> next token) is what
 * bounds a decoder that handles one sequence per step.  Here it is cut twice:
 *
 *  parse   Every lane decodes the bytes at compressed position ip + lane AS IF
 *          a token started there (one or two unaligned 4-byte reads).  Which
 *          of the 32 positions really are token starts follows from pointer
 *          doubling over the lanes' "next token" distances (4 shuffle rounds:
 *          a hop is at least 3 bytes, so a 32-byte window holds at most 11
 *          tokens).  The real sequences are appended to a descriptor ring in
 *          shared memory {literal position, literal length, offset, match
 *          length}; no data moves yet.
 *  copy    When 32 descriptors are queued (or the regular stretch ends) one
 *          lane takes one sequence: a prefix sum over the lengths gives every
 *          sequence its output position.  The wave's literal runs and matches
 *          become SEGMENTS of its output bytes: a bitmap marks the first byte
 *          of each, a table holds what a byte of the segment is copied from
 *          (the compressed stream at a fixed distance, or `offset` bytes
 *          back).  Then the bytes go out 32 at a time in output order, one
 *          byte per lane: rank of the byte in the bitmap -> its segment ->
 *          its source.  Every lane always has a byte to move, whatever the
 *          sequence lengths are.  A source that lies inside the same 32 bytes
 *          is followed further back (an overlapping match by whole periods)
 *          until it reaches a literal or bytes of an earlier step, so
 *          overlapping matches and chains of short matches need no special
 *          path; offset 0 cannot occur here (it is irregular).
 */
#ifndef LZ4E_DECOMPRESS_WAVE_CUH
#define LZ4E_DECOMPRESS_WAVE_CUH

#include "lz4e_decompress_group.cuh"

#define LZ4E_W_QCAP 64u	  /* descriptor ring entries (power of two, >= 32 + 11) */
#define LZ4E_W_LITMAX 256u /* literal runs of a regular sequence: below this */
#define LZ4E_W_TMAX 2048u  /* output bytes of one wave */
#define LZ4E_W_BIAS 4096u  /* keeps "compressed - output position" positive */
#define LZ4E_W_RING 4096u  /* most recent output bytes kept in shared memory */

/* per-warp shared memory */
struct lz4e_wave_smem {
	lz4e_omap map;		   /* destination segment map of the block */
	uint32_t qx[LZ4E_W_QCAP]; /* literal position (24) | literal length (8) */
	uint32_t qy[LZ4E_W_QCAP]; /* offset (16) | match length (16) */
	uint32_t bm[LZ4E_W_TMAX / 32u]; /* first bytes of the wave's segments */
	uint32_t seg[64];		 /* one word per segment */
	/*
	 * The last LZ4E_W_RING output bytes of the regular stretch, at index
	 * (output position mod LZ4E_W_RING).  A match byte whose source is that
	 * close is read from here instead of from the destination in global
	 * memory: just-written output is not in L1 (stores write through and
	 * invalidate), so every such read was an L2 round trip — the decoder's
	 * largest stall — and, for scattered destinations, a segment lookup.
	 */
	uint8_t ring[LZ4E_W_RING];
};

LZ4E_DEV uint64_t lz4e_w_shfl_ptr(const void *p, int src)
{
	return lz4e_shfl64((uint64_t)(uintptr_t)p, src);
}

/* address holding output byte `pos` (already written) of the current block */
LZ4E_DEV const uint8_t *lz4e_w_outaddr(const lz4e_gstate<32, false> &g, const lz4e_omap *m,
				       uint32_t pos)
{
	const uint8_t *src = g.ob + pos;

	if (g.nseg > 1u && pos < g.o_lo) {
		if (pos >= g.p_lo) {
			src = g.pb + pos;
		} else {
			uint32_t k = g.oseg - 1u;

			while (m->start[k] > pos)
				k--;
			src = (const uint8_t *)(uintptr_t)m->bias[k] + pos;
		}
	}
	LZ4E_ASSUME_GLOBAL(src);
	return src;
}

/*
 * Regular sequences of the current block until the next one is not regular.
 * On return g.ip / g.op name that sequence; everything before it is written.
 */
LZ4E_DEV void lz4e_w_hot(lz4e_gstate<32, false> &g, lz4e_wave_smem *ws, uint32_t lane,
			  uint32_t serial_min)
{
	const uint8_t *inb = g.inb;
	uint8_t *ob = g.ob, *pb = g.pb;
	const lz4e_omap *m = &ws->map;
	const uint32_t o_lo = g.o_lo, p_lo = g.p_lo, nseg = g.nseg, oseg = g.oseg;
	const uint32_t ihi = lz4e_min(g.in_hi, g.iend);
	const uint32_t lim = g.oend >= 12u ? lz4e_min(g.o_hi, g.oend - 12u) : 0u;
	const uint32_t lt = (1u << lane) - 1u;
	uint32_t ip = g.ip, op = g.op;
	uint32_t qh = 0, qn = 0;
	bool stop = false;
	/* the ring holds output positions [max(ring_lo, front - RING), front):
	 * what was produced before this call (by-the-book sequences, the previous
	 * block) is not in it */
	const uint32_t ring_lo = op;
	uint8_t *ring = ws->ring;

	LZ4E_ASSUME_GLOBAL(inb);
	LZ4E_ASSUME_GLOBAL(ob);
	LZ4E_ASSUME_GLOBAL(pb);
	for (;;) {
		/* ---- parse: windows of 32 compressed bytes ---- */
		while (qn < 32u && !stop) {
			const uint32_t p = ip + lane;
			const bool v = p + 8u <= ihi;
			const uint32_t w = v ? lz4e_ld32u(inb + p) : 0u;
			const uint32_t l0 = (w >> 4) & 15u, ml4 = w & 15u;
			const uint32_t e1 = (w >> 8) & 0xffu;
			const bool lext = l0 == 15u;
			const uint32_t lit = l0 + (lext ? e1 : 0u);
			const uint32_t lp = p + 1u + (lext ? 1u : 0u); /* first literal */
			const uint32_t fp = lp + lit;		       /* offset field */
			const bool ok = v && !(lext && e1 == 255u) && fp + 15u <= ihi;
			uint32_t ow = w >> 8;
			const bool far = ok && fp != p + 1u;

			if (lz4e_ballot(LZ4E_FULL, far)) {
				if (far)
					ow = lz4e_ld32u(inb + fp);
			}
			const uint32_t off = ow & 0xffffu;
			const uint32_t e2 = (ow >> 16) & 0xffu;
			const bool mext = ml4 == 15u;
			const uint32_t mlen = ml4 + 4u + (mext ? e2 : 0u);
			const bool simple = ok && !(mext && e2 == 255u) && off != 0u &&
					    lit < LZ4E_W_LITMAX;
			const uint32_t nxt = fp + 2u + (mext ? 1u : 0u) - ip; /* >= lane + 3 */
			/* which lanes are token starts: reachable from lane 0 */
			uint32_t hop = simple ? lz4e_min(nxt, 32u) : 32u;
			uint32_t mk = 1u << lane;

#pragma unroll
			for (int r = 0; r < 4; r++) {
				const uint32_t m2 = lz4e_shfl(LZ4E_FULL, mk, (int)(hop & 31u));
				const uint32_t h2 = lz4e_shfl(LZ4E_FULL, hop, (int)(hop & 31u));

				if (hop < 32u) {
					mk |= m2;
					hop = h2;
				}
			}
			const uint32_t real = lz4e_shfl(LZ4E_FULL, mk, 0);
			const uint32_t irr = lz4e_ballot(LZ4E_FULL, !simple) & real;
			/* real sequences before the first irregular one */
			const uint32_t acc = irr ? (real & ((irr & (0u - irr)) - 1u)) : real;

			if (irr)
				stop = true;
			if (acc == 0u)
				break;
			if ((acc >> lane) & 1u) {
				const uint32_t slot =
					(qh + qn + (uint32_t)lz4e_popc(acc & lt)) & (LZ4E_W_QCAP - 1u);

				ws->qx[slot] = lp | (lit << 24);
				ws->qy[slot] = off | (mlen << 16);
			}
			ip += lz4e_shfl(LZ4E_FULL, nxt, 31 - lz4e_clz(acc));
			qn += (uint32_t)lz4e_popc(acc);
		}
		if (qn == 0u)
			break; /* stop: the sequence at ip is irregular */
		lz4e_syncwarp(LZ4E_FULL);

		/* ---- copy ---- */
		const uint32_t n = lz4e_min(qn, 32u);
		uint32_t dx = 0, dy = 0;

		if (lane < n) {
			const uint32_t slot = (qh + lane) & (LZ4E_W_QCAP - 1u);

			dx = ws->qx[slot];
			dy = ws->qy[slot];
		}
		const uint32_t lp = dx & 0xffffffu, lit = dx >> 24;
		const uint32_t off = dy & 0xffffu, mlen = dy >> 16;
		const uint32_t tot = lit + mlen;
		uint32_t incl = tot;

#pragma unroll
		for (uint32_t d = 1; d < 32u; d <<= 1) {
			const uint32_t t = lz4e_shfl(LZ4E_FULL, incl, (int)((lane - d) & 31u));

			if (lane >= d)
				incl += t;
		}
		const uint32_t r = incl - tot; /* first output byte, relative to op */
		const uint32_t mo = op + r + lit; /* match destination */
		/* sequences of this wave: as many as fit LZ4E_W_TMAX output bytes, up
		 * to the first one that is not regular after all (offset beyond the
		 * output produced so far, or too close to the end of the block or of
		 * the destination segment) */
		const uint32_t fitm = lz4e_ballot(LZ4E_FULL, lane < n && incl <= LZ4E_W_TMAX);
		const uint32_t badm =
			lz4e_ballot(LZ4E_FULL, lane < n && (off > mo || mo + mlen > lim)) & fitm;
		const uint32_t na = badm ? (uint32_t)lz4e_ffs(badm) - 1u : (uint32_t)lz4e_popc(fitm);
		const bool act = lane < na;
		const uint32_t T = na ? lz4e_shfl(LZ4E_FULL, incl, (int)(na - 1u)) : 0u;

		/*
		 * Long sequences (serial_min output bytes or more on average): one
		 * sequence at a time, all lanes on its bytes, 32 per pass.  A pass
		 * costs about a third of a byte-parallel step below and needs no
		 * segment table; the byte-parallel form keeps every lane busy
		 * whatever the lengths are and wins on short sequences (text:
		 * 8 bytes per sequence).  Sources come from the ring when they are
		 * close, and a match with an offset below 32 is its first period
		 * repeated, so no pass reads what the same pass writes.
		 */
		if (na && serial_min && T >= serial_min * na) {
			for (uint32_t q = 0; q < na; q++) {
				const uint32_t sx = lz4e_shfl(LZ4E_FULL, dx, (int)q);
				const uint32_t sy = lz4e_shfl(LZ4E_FULL, dy, (int)q);
				const uint32_t o0 = op + lz4e_shfl(LZ4E_FULL, r, (int)q);
				const uint32_t s_lp = sx & 0xffffffu, s_lit = sx >> 24;
				const uint32_t s_off = sy & 0xffffu, s_ml = sy >> 16;
				const uint32_t m0 = o0 + s_lit;

				if (s_lit) {
					for (uint32_t i = lane; i < s_lit; i += 32u) {
						const uint8_t v = inb[s_lp + i];

						ob[o0 + i] = v;
						ring[(o0 + i) & (LZ4E_W_RING - 1u)] = v;
					}
					lz4e_syncwarp(LZ4E_FULL);
				}
				const float rc = lz4e_rcpf((float)s_off);

				for (uint32_t c = 0; c < s_ml; c += 32u) {
					const uint32_t i = c + lane;
					/* offset below 32: byte i of the first period */
					const uint32_t k =
						s_off < 32u
							? i - s_off * (uint32_t)(((float)i + 0.5f) * rc)
							: i;
					const uint32_t pos = m0 - s_off + k;
					const uint32_t front = m0 + c;

					if (i < s_ml) {
						uint8_t v;

						if (pos >= ring_lo && front - pos <= LZ4E_W_RING - 32u)
							v = ring[pos & (LZ4E_W_RING - 1u)];
						else
							v = *lz4e_w_outaddr(g, m, pos);
						ob[m0 + i] = v;
						ring[(m0 + i) & (LZ4E_W_RING - 1u)] = v;
					}
					lz4e_syncwarp(LZ4E_FULL);
				}
			}
		} else {
		/*
		 * Segment table of the wave: every literal run and every match is
		 * one segment of the wave's T output bytes; bm has a bit for the
		 * first byte of each, seg[] one word per segment in output order:
		 * literal run: bit 31 | (compressed position - output position),
		 * biased; match: the offset.
		 */
		{
			const uint32_t zl = lz4e_ballot(LZ4E_FULL, act && lit == 0u);
			const uint32_t il = 2u * lane - (uint32_t)lz4e_popc(zl & lt);
			const uint32_t nw = (T + 31u) >> 5;

			if (lane < nw)
				ws->bm[lane] = 0;
			if (lane + 32u < nw)
				ws->bm[lane + 32u] = 0;
			lz4e_syncwarp(LZ4E_FULL);
			if (act) {
				const uint32_t rm = r + lit;

				if (lit) {
					lz4e_atomic_or(&ws->bm[r >> 5], 1u << (r & 31u));
					ws->seg[il] = 0x80000000u | (lp + LZ4E_W_BIAS - r);
				}
				lz4e_atomic_or(&ws->bm[rm >> 5], 1u << (rm & 31u));
				ws->seg[il + (lit ? 1u : 0u)] = off;
			}
			lz4e_syncwarp(LZ4E_FULL);
		}
		/*
		 * The wave's bytes, 32 at a time in output order: lane l moves byte
		 * c0 + l.  A match byte whose source lies inside the same 32 bytes
		 * follows the chain back (whole periods of an overlapping match at
		 * once) until it reaches a literal or a byte written before this
		 * step.
		 */
		{
			const uint32_t le = lt | (1u << lane);
			uint32_t base = 0; /* segments that start before this step */

			for (uint32_t c0 = 0; c0 < T; c0 += 32u) {
				const uint32_t w = ws->bm[c0 >> 5];
				const uint32_t j = c0 + lane;
				const bool on = j < T;
				const uint32_t e = ws->seg[base + (uint32_t)lz4e_popc(w & le) - 1u];
				bool isl = (e >> 31) != 0u;
				/* literal: compressed position; match: output position
				 * relative to op (negative: before this wave) */
				int32_t sp = isl ? (int32_t)(j + (e & 0x7fffffffu) - LZ4E_W_BIAS)
						 : (int32_t)j - (int32_t)e;

				for (;;) {
					const bool in = on && !isl && sp >= (int32_t)c0;

					if (lz4e_ballot(LZ4E_FULL, in) == 0)
						break;
					if (in) {
						const uint32_t q = (uint32_t)sp - c0;
						const uint32_t wq = w & (0xffffffffu >> (31u - q));
						const uint32_t e2 = ws->seg[base + (uint32_t)lz4e_popc(wq) - 1u];

						if (e2 >> 31) {
							isl = true;
							sp += (int32_t)((e2 & 0x7fffffffu) - LZ4E_W_BIAS);
						} else {
							/* first byte of that match inside this
							 * step; go back whole periods past it */
							const uint32_t lb = wq ? 31u - (uint32_t)lz4e_clz(wq) : 0u;
							const uint32_t k =
								(uint32_t)(((float)(q - lb) + 0.5f) *
									   lz4e_rcpf((float)e2)) + 1u;

							sp -= (int32_t)(k * e2);
						}
					}
				}
				const uint32_t pos = op + (uint32_t)sp; /* match: linear output position */
				/* this step overwrites ring positions front - RING ..
				 * front - RING + 31: sources that old go to memory */
				const uint32_t front = op + c0;
				const bool near = on && !isl && pos >= ring_lo &&
						  front - pos <= LZ4E_W_RING - 32u;
				const bool far = on && !near;
				uint32_t v = 0;

				if (near)
					v = ring[pos & (LZ4E_W_RING - 1u)];
				if (lz4e_ballot(LZ4E_FULL, far)) {
					const uint8_t *src = isl ? inb + (uint32_t)sp : ob + pos;

					if (nseg > 1u) {
						const bool back = far && !isl && pos < o_lo;

						if (lz4e_ballot(LZ4E_FULL, back)) {
							if (back) {
								if (pos >= p_lo) {
									src = pb + pos;
								} else {
									uint32_t k = oseg - 1u;

									while (m->start[k] > pos)
										k--;
									src = (const uint8_t *)(uintptr_t)m->bias[k] + pos;
								}
							}
						}
					}
					LZ4E_ASSUME_GLOBAL(src);
					if (far)
						v = *src;
				}
				if (on) {
					ob[op + j] = (uint8_t)v;
					ring[(op + j) & (LZ4E_W_RING - 1u)] = (uint8_t)v;
				}
				base += (uint32_t)lz4e_popc(w);
				lz4e_syncwarp(LZ4E_FULL);
			}
		}
		}
		if (na) {
			op += T;
			qh = (qh + na) & (LZ4E_W_QCAP - 1u);
			qn -= na;
		}
		if (badm) {
			/* the sequence of lane `na` goes by the book; what the parse
			 * queued behind it is dropped */
			const uint32_t tp = lp - 1u - (lit >= 15u ? 1u : 0u);

			ip = lz4e_shfl(LZ4E_FULL, tp, (int)na);
			break;
		}
	}
	g.ip = ip;
	g.op = op;
	g.irregular = true;
}

/*
 * ws: this warp's shared memory; slot: this warp's scratch slot (used only for
 * destinations of more than LZ4E_G_MAXSEG segments).
 */
LZ4E_DEV void lz4e_decompress_wave(const lz4e_decompress_args &a, lz4e_wave_smem *ws,
				   uint32_t slot, uint32_t lane)
{
	lz4e_gstate<32, false> g;

	g.inb = (const uint8_t *)a.src_bvecs;
	g.sv = a.src_bvecs;
	g.ob = g.pb = (uint8_t *)a.rets;
	g.in_hi = g.iend = g.o_lo = g.o_hi = g.p_lo = g.oend = 0;
	g.oseg = 0;
	g.nseg = 1;
	g.ip = g.op = g.b = 0;
	g.active = g.done = g.scatter = g.irregular = false;

	for (;;) {
		if (!g.active) {
			lz4e_g_open(a, g, &ws->map, (uint8_t *)nullptr, slot, lane, LZ4E_FULL, 0);
			if (g.done)
				return;
			if (!g.active)
				continue;
		} else if (g.irregular) {
			lz4e_g_slow(a, g, &ws->map, lane, LZ4E_FULL);
			if (!g.active)
				continue;
		}
		lz4e_syncwarp(LZ4E_FULL);
		lz4e_w_hot(g, ws, lane, a.wave_serial);
	}
}

#endif /* LZ4E_DECOMPRESS_WAVE_CUH */
