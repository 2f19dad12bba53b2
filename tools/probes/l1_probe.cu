// l1_probe.cu — does a global store keep / update / invalidate the storing SM's L1 line?
// Decides how the LZ4E decoder should read back bytes it has just written (match copies).
// Every scenario is a dependent chain over N fresh 128-byte lines, timed as a whole;
// the table printed is cycles per line.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l1_probe l1_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t ld8(const uint8_t *p)
{
	uint32_t v;
	asm volatile("ld.global.u8 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st8(uint8_t *p, uint32_t v)
{
	asm volatile("st.global.u8 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st128(uint8_t *p, uint32_t v)
{
	asm volatile("st.global.v4.u32 [%0], {%1,%1,%1,%1};" ::"l"(p), "r"(v) : "memory");
}

#define N 256

// mode: sequence of operations per line; every load's address depends on the previous load
__global__ void probe(uint8_t *buf, long long *out, int mode)
{
	const int lane = threadIdx.x;
	uint32_t v = 0;
	long long t0 = 0, t1 = 0;

	if (mode < 8) {
		if (lane != 0)
			return;
		t0 = clock64();
		for (int i = 0; i < N; i++) {
			uint8_t *p = buf + (size_t)i * 128 + (v & 64); /* v is always 1: & 64 == 0 */
			switch (mode) {
			case 0: /* ld (miss) */
				v = ld8(p);
				break;
			case 1: /* ld (miss), ld (hit) */
				v = ld8(p);
				v = ld8(p + 1 + (v & 64));
				break;
			case 2: /* ld, st same sector, ld */
				v = ld8(p);
				st8(p + 1 + (v & 64), 1);
				v = ld8(p + 1 + (v & 64));
				break;
			case 3: /* ld, st same sector, ld, ld */
				v = ld8(p);
				st8(p + 1 + (v & 64), 1);
				v = ld8(p + 1 + (v & 64));
				v = ld8(p + 2 + (v & 64));
				break;
			case 4: /* st (line never loaded), ld */
				st8(p, 1);
				v = ld8(p + (v & 64));
				break;
			case 5: /* st, ld, ld */
				st8(p, 1);
				v = ld8(p + (v & 64));
				v = ld8(p + 1 + (v & 64));
				break;
			case 6: /* st, ld, st, ld */
				st8(p, 1);
				v = ld8(p + (v & 64));
				st8(p + 1 + (v & 64), 1);
				v = ld8(p + 1 + (v & 64));
				break;
			case 7: /* st.v4, ld, st.v4 other sector, ld it */
				st128(p, 0x01010101u);
				v = ld8(p + (v & 64));
				st128(p + 32 + (v & 64), 0x01010101u);
				v = ld8(p + 33 + (v & 64));
				break;
			}
		}
		t1 = clock64();
	} else {
		/* cross-lane: lane 1 stores, lane 0 loads, __syncwarp between (the decoder's pattern) */
		t0 = clock64();
		for (int i = 0; i < N; i++) {
			uint8_t *p = buf + (size_t)i * 128 + (v & 64);
			if (mode == 8) { /* lane0 ld; lane1 st; lane0 ld */
				if (lane == 0)
					v = ld8(p);
				v = __shfl_sync(0xffffffffu, v, 0);
				if (lane == 1)
					st8(p + 1 + (v & 64), 1);
				__syncwarp();
				if (lane == 0)
					v = ld8(p + 1 + (v & 64));
				v = __shfl_sync(0xffffffffu, v, 0);
			} else if (mode == 9) { /* same without the store (baseline) */
				if (lane == 0)
					v = ld8(p);
				v = __shfl_sync(0xffffffffu, v, 0);
				__syncwarp();
				if (lane == 0)
					v = ld8(p + 1 + (v & 64));
				v = __shfl_sync(0xffffffffu, v, 0);
			} else if (mode == 10) { /* 32 lanes store a 32-byte row, then all load row shifted by 8 */
				st8(p + lane, 1);
				__syncwarp();
				v = ld8(p + ((lane + 8) & 31) + (v & 64));
				v = __shfl_sync(0xffffffffu, v, 0);
			} else { /* mode 11: line pre-loaded by all, then st row, then ld */
				v = ld8(p + lane);
				st8(p + lane + (v & 64), 1);
				__syncwarp();
				v = ld8(p + ((lane + 8) & 31) + (v & 64));
				v = __shfl_sync(0xffffffffu, v, 0);
			}
		}
		t1 = clock64();
	}
	if (lane == 0) {
		out[mode * 2] = t1 - t0;
		out[mode * 2 + 1] = v;
	}
}

int main()
{
	uint8_t *buf;
	long long *out, h[32];
	const char *names[12] = {
		"0  ld(miss)",
		"1  ld(miss) ld",
		"2  ld(miss) st ld",
		"3  ld(miss) st ld ld",
		"4  st ld",
		"5  st ld ld",
		"6  st ld st ld",
		"7  st.v4 ld st.v4(other sector) ld",
		"8  lane0 ld(miss); lane1 st; syncwarp; lane0 ld   (+2 shfl)",
		"9  lane0 ld(miss); syncwarp; lane0 ld             (+2 shfl)",
		"10 32 lanes st row; syncwarp; ld row              (+1 shfl)",
		"11 32 lanes ld(miss) row; st row; syncwarp; ld    (+1 shfl)",
	};
	cudaMalloc(&buf, 16 << 20);
	cudaMalloc(&out, sizeof(h));
	cudaMemset(out, 0, sizeof(h));
	for (int m = 0; m < 12; m++) {
		cudaMemset(buf, 1, 16 << 20); /* L2-resident, not in any L1 (new launch) */
		probe<<<1, 32>>>(buf + (size_t)m * (1 << 20), out, m);
		cudaDeviceSynchronize();
	}
	cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
	for (int m = 0; m < 12; m++)
		printf("%-64s %7.1f cycles/line\n", names[m], (double)h[m * 2] / N);
	printf("cuda status: %s\n", cudaGetErrorString(cudaGetLastError()));
	return 0;
}
