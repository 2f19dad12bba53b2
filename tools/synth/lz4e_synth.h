/*
 * Synthetic block contents for the benchmark configs of BASELINE.json
 * (SURVEY.md section 8d, config 2): content class = block_index mod 4,
 * per-block seed = block index.  One definition shared by the host library
 * (tests, CPU baseline inputs) and the device fill kernel (bench.py), so both
 * sides compress exactly the same bytes.
 *
 *   class 0  word text     LCG-picked words from a 12-word list + ' '
 *   class 1  random bytes  incompressible
 *   class 2  sparse page   zeros with a non-zero byte every ~97 bytes
 *   class 3  mixed         short literal runs, byte runs and back-references
 *                          at offsets 1..60000 (short and long matches,
 *                          overlapping copies)
 */
#ifndef LZ4E_SYNTH_H
#define LZ4E_SYNTH_H

#include <stdint.h>

#ifdef __CUDACC__
#define LZ4E_HD __host__ __device__
#else
#define LZ4E_HD
#endif

LZ4E_HD static inline uint32_t lz4e_lcg(uint32_t *s)
{
	*s = *s * 1664525u + 1013904223u;
	return *s;
}

LZ4E_HD static inline void lz4e_synth_block(uint8_t *buf, uint32_t n,
					    uint32_t cls, uint32_t seed)
{
	/* "lorem ipsum dolor sit amet consectetur adipiscing elit sed do
	 * eiusmod tempor", fixed-width table so it lives in registers/const */
	const char words[12][12] = { "lorem",	    "ipsum", "dolor",
				     "sit",	    "amet",  "consectetur",
				     "adipiscing",  "elit",  "sed",
				     "do",	    "eiusmod", "tempor" };
	uint32_t s = seed, p = 0;

	switch (cls & 3u) {
	case 0:
		while (p < n) {
			const char *w = words[(lz4e_lcg(&s) >> 16) % 12u];
			uint32_t k = 0;

			while (w[k] && p < n)
				buf[p++] = (uint8_t)w[k++];
			if (p < n)
				buf[p++] = ' ';
		}
		break;
	case 1:
		while (p < n)
			buf[p++] = (uint8_t)(lz4e_lcg(&s) >> 24);
		break;
	case 2:
		for (p = 0; p < n; p++)
			buf[p] = 0;
		p = (lz4e_lcg(&s) >> 16) % 193u;
		while (p < n) {
			buf[p] = (uint8_t)((lz4e_lcg(&s) >> 24) | 1u);
			p += 1u + (lz4e_lcg(&s) >> 16) % 193u;
		}
		break;
	default:
		while (p < n) {
			uint32_t r = lz4e_lcg(&s);
			uint32_t kind = (r >> 28) & 3u;

			if (kind == 0 || p == 0) {
				uint32_t len = 1u + ((r >> 8) & 15u);

				while (len-- && p < n)
					buf[p++] = (uint8_t)(lz4e_lcg(&s) >> 24);
			} else if (kind == 1) {
				uint32_t len = 4u + ((r >> 8) & 63u);
				uint8_t v = (uint8_t)(r >> 16);

				while (len-- && p < n)
					buf[p++] = v;
			} else {
				uint32_t r2 = lz4e_lcg(&s);
				uint32_t span = p < 60000u ? p : 60000u;
				uint32_t off = 1u + (r2 >> 8) % span;
				uint32_t len = 4u + (r >> 8) % 60u;

				if ((r2 & 7u) == 0)
					len += (r >> 16) & 511u;
				while (len-- && p < n) {
					buf[p] = buf[p - off];
					p++;
				}
			}
		}
		break;
	}
}

#endif /* LZ4E_SYNTH_H */
