// tools/lds_probe.cu -- MEASUREMENT INFRASTRUCTURE: shared-memory load throughput for the
// profile access pattern of sab_wave.cuh (every lane reads entry [column][its own letter],
// the column being an immediate offset) with 32-, 64- and 128-bit loads and L distinct letters.
// Answers: does one LDS.64/.128 whose lanes touch <= 128 B cost one wavefront or 2/4?
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

template <int W> __device__ __forceinline__ uint32_t lds(uint32_t addr)
{
	uint32_t a, b, c, d;
	if (W == 4)
		asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(addr));
	else if (W == 8)
		asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
	else
		asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
	return a;
}

// table layout: [column][letter] entries of W bytes, column stride = (L*W + pad) bytes
template <int W> __global__ void __launch_bounds__(128) probe(int iters, int L, int stride, uint32_t seed, uint32_t *out)
{
	extern __shared__ __align__(16) unsigned char smem[];
	const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
	uint32_t s = seed * 2654435761u + threadIdx.x * 40503u + blockIdx.x;
	uint32_t pq[8];
	for (int k = 0; k < 8; k++) {
		s = s * 1664525u + 1013904223u;
		pq[k] = base + ((s >> 16) % L) * W;
	}
	uint32_t acc[4] = { 0, 0, 0, 0 };
	for (int it = 0; it < iters; it++) {
#pragma unroll
		for (int c = 0; c < 8; c++)
#pragma unroll
			for (int k = 0; k < 8; k++)
				acc[k & 3] += lds<W>(pq[k] + c * stride);
	}
	if ((acc[0] ^ acc[1] ^ acc[2] ^ acc[3]) == 0x12345u)
		out[threadIdx.x] = pq[0];
}

static uint32_t *dout;
template <int W> static void run(int L, int pad, int sms, double mhz)
{
	const int stride = L * W + pad;
	const int iters = 4096, blocks_per_sm = 8;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	for (int rep = 0; rep < 4; rep++) {
		cudaEventRecord(e0);
		probe<W><<<sms * blocks_per_sm, 128, 8 * stride + 64>>>(iters, L, stride, 17u + rep, dout);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (rep && ms < best)
			best = ms;
	}
	const double inst_per_sm = (double)blocks_per_sm * 4 * iters * 64;
	const double clk = best * 1e-3 * mhz * 1e6;
	printf("LDS.%-3d letters %2d stride %4d B: %.3f warp-loads/clk/SM, %.1f useful B/clk/SM (%s)\n", W * 8, L, stride,
	       inst_per_sm / clk, inst_per_sm / clk * 32 * W, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
	cudaDeviceProp prop;
	cudaGetDeviceProperties(&prop, 0);
	int khz = 0;
	cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
	const double mhz = khz / 1e3;
	const int sms = prop.multiProcessorCount;
	cudaMalloc(&dout, 4096);
	const int Ls[] = { 1, 4, 8, 16, 20, 25 };
	for (int L : Ls)
		run<4>(L, L == 25 ? 0 : 4, sms, mhz);
	for (int L : Ls)
		run<8>(L, 8, sms, mhz);
	for (int L : Ls)
		run<16>(L, 16, sms, mhz);
	return 0;
}
