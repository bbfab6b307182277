// tools/cell_peak.cu -- MEASUREMENT INFRASTRUCTURE: ceiling of the DP cell code shape.
//
// Runs the 8-row x K-column register-strip block of sab_wave.cuh (same cell<> function, same
// anti-diagonal issue order, same shared-memory profile lookup) in a bare loop: no work
// queue, no strip boundaries, no profile staging, optionally no per-step barrier.  What it
// reports (clk per packed cell pair per SM sub-partition) is the most the real kernel could
// reach with this instruction mix and occupancy; the gap between the two is what the
// pipeline plumbing costs.  Not part of the product path.
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I sequencealigner_b200/csrc
//        -I include -o tools/cell_peak tools/cell_peak.cu ; tools/cell_peak
// With -DCELL_PEAK_LIB only sab_cell_peak() is built (bench.py, libsab_pipepeak.so).
#include "sab_wave.cuh"

#include <cstdio>
#include <vector>

using namespace sab;

template <int METHOD, int K, int MINB, bool LDS_ON, bool BAR, int NR = 8>
__global__ void __launch_bounds__(128, MINB) micro(uint32_t *out, int iters, int gA, int gE, uint32_t one,
						   const uint32_t *letters)
{
	using O = VOps<true>;
	extern __shared__ uint32_t dyn_pad[];
	__shared__ uint32_t prof[4][K][kProfStride];
	__shared__ uint32_t lets[64][32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	for (int i = threadIdx.x; i < 4 * K * kProfStride; i += 128)
		(&prof[0][0][0])[i] = O::delta2((i * 7) % 9 - 4 - 2 * gE, (i * 5) % 11 - 4 - 2 * gE);
	for (int i = threadIdx.x; i < 64 * 32; i += 128)
		(&lets[0][0])[i] = letters[i];
	__syncthreads();
	CellK<METHOD, true> ck;
	ck.one = one;
#ifdef SAB_IMM_LG
	ck.lG = METHOD == M_GA ? 0xfff7fff7u : 0xffffffffu;
#else
	ck.lG = O::lane(METHOD == M_GA ? gA - gE : gE);
#endif
	ck.dE = O::delta(gE);
	ck.dA = O::delta(gA);
	ck.ZERO = O::bcast(0);
	ck.gE = gE;
	uint32_t V[K], Y[K];
#pragma unroll
	for (int c = 0; c < K; c++) {
		V[c] = O::bcast(c + lane);
		Y[c] = O::bcast(0);
	}
	uint32_t diag = O::bcast(0), bestA = ck.ZERO, bestB = ck.ZERO;
	uint2 carry[NR];
#pragma unroll
	for (int k = 0; k < NR; k++)
		carry[k] = make_uint2(O::bcast(k), O::bcast(0));
	const uint32_t *prof_s = prof[warp][0];
	for (int it = 0; it < iters; it++) {
		const uint32_t w0 = lets[(it * 2) & 63][lane], w1 = lets[(it * 2 + 1) & 63][lane];
		const uint32_t w2 = NR > 8 ? lets[(it * 2 + 2) & 63][lane] : w0, w3 = NR > 8 ? lets[(it * 2 + 3) & 63][lane] : w1;
		uint32_t ml[NR], x[NR], dg[NR], mprev[NR];
		const uint32_t *pq[NR];
#pragma unroll
		for (int k = 0; k < NR; k++) {
			ml[k] = carry[k].x;
			x[k] = carry[k].y;
			dg[k] = k ? carry[k - 1].x : diag;
			mprev[k] = 0;
			const uint32_t wk = (k & 8) ? ((k & 4) ? w3 : w2) : ((k & 4) ? w1 : w0);
			pq[k] = reinterpret_cast<const uint32_t *>(reinterpret_cast<const char *>(prof_s) +
								   ((wk >> (8 * (k & 3))) & 0xffu));
		}
		diag = carry[NR - 1].x;
#pragma unroll
		for (int tt = 0; tt < K + NR - 1; tt++) {
#pragma unroll
			for (int k = 0; k < NR; k++) {
				const int c = tt - k;
				if (c < 0 || c >= K)
					continue;
				const uint32_t s = LDS_ON ? pq[k][c * kProfStride] : (uint32_t)(size_t)pq[k];
				const uint32_t m = cell<METHOD, true>(ck, V[c], Y[c], ml[k], x[k], dg[k], s);
				if (METHOD == M_SW) {
					if (c & 1) {
						if (k & 1)
							bestB = O::max3(bestB, mprev[k], m);
						else
							bestA = O::max3(bestA, mprev[k], m);
					} else {
						mprev[k] = m;
					}
				}
			}
		}
#pragma unroll
		for (int k = 0; k < NR; k++)
			carry[k] = make_uint2(ml[k], x[k]);
		if (BAR)
			__syncthreads();
	}
	uint32_t acc = bestA ^ bestB ^ diag;
#pragma unroll
	for (int c = 0; c < K; c++)
		acc ^= V[c] ^ Y[c];
	out[blockIdx.x * 128 + threadIdx.x] = acc;
}

template <int METHOD, int K, int MINB, bool LDS_ON, bool BAR, int NR = 8>
static double run(const char *name, uint32_t *out, const uint32_t *letters, int sms, double mhz, int occ_cap = 0,
		  double *clk_out = nullptr)
{
	const int iters = 20000 / (K / 16) * 8 / NR;
	auto kern = micro<METHOD, K, MINB, LDS_ON, BAR, NR>;
	cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
	int dyn = 0;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	int occ = 0;
	cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, 0);
	while (occ_cap && occ > occ_cap) {
		dyn += 2048;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, dyn);
	}
	const int blocks = sms * occ;
	for (int rep = 0; rep < 4; rep++) {
		cudaEventRecord(e0);
		kern<<<blocks, 128, dyn>>>(out, iters, -10, -1, 1u, letters);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (rep && ms < best)
			best = ms;
	}
	const double pairs_per_smsp = (double)occ * iters * (double)NR * K; // one warp per SMSP per block
	const double clk = best * 1e-3 * mhz * 1e6 / pairs_per_smsp;
	const double tcups = (double)blocks * 4 * iters * (double)NR * K * 64 / (best * 1e-3) / 1e12;
	if (name)
		printf("%-34s blocks/SM %d  %.2f ms  %.2f clk/cell-pair/SMSP (at %.0f MHz)  %.2f TCUPS\n", name,
		       occ, best, clk, mhz, tcups);
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) {
		if (name)
			printf("  CUDA error: %s\n", cudaGetErrorString(e));
		return -1.0;
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	if (clk_out)
		*clk_out = clk;
	return tcups;
}

// Scheduler probe: ILP independent GA-shaped cells per thread (3 DPX + IMAD [+ LDS]), no strip
// structure at all.  Shows what the SM issues for this instruction MIX at a given occupancy.
template <int MODE, int ILP> __global__ void __launch_bounds__(128) mixk(uint32_t *out, int iters, uint32_t one, uint32_t seed)
{
	extern __shared__ uint32_t dyn_pad[];
	__shared__ uint32_t sm[32 * 64];
	for (int i = threadIdx.x; i < 32 * 64; i += 128)
		sm[i] = (i * 3) & 0x00070007u;
	__syncthreads();
	const volatile uint32_t *ps = sm + (threadIdx.x & 31);
	uint32_t v[ILP], w[ILP], x[ILP];
	for (int k = 0; k < ILP; k++) {
		v[k] = seed * (k + 1) + threadIdx.x;
		w[k] = seed ^ (k * 77 + threadIdx.x);
		x[k] = seed + k;
	}
	const uint32_t c1 = seed | 0x00010001u;
	uint32_t acc2[4] = { seed, seed + 1, seed + 2, seed + 3 };
	float facc[4] = { 1.f, 2.f, 3.f, 4.f };
	const float fone = __uint_as_float(seed);
	uint32_t lacc = 0;
	uint32_t hacc[8];
	for (int k = 0; k < 8; k++)
		hacc[k] = seed + k * 0x00010001u;
	const uint32_t pb = (uint32_t)__cvta_generic_to_shared(sm) + ((seed * (threadIdx.x + 7)) % 20) * 4;
	for (int it = 0; it < iters; it++) {
#pragma unroll
		for (int rep = 0; rep < 4; rep++) {
#pragma unroll
			for (int k = 0; k < ILP; k++) {
				if (MODE >= 3) { // independent streams: pure throughput of the mix
					if (MODE == 3 || MODE == 5 || MODE == 6) {
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
					}
					if (MODE == 4)
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
					if (MODE == 4 || MODE == 5) {
						const uint32_t s = ps[((rep * ILP + k) & 63) * 32];
						acc2[k & 3] = imad_add(acc2[k & 3], one, s);
					}
					if (MODE == 6)
						acc2[k & 3] = imad_add(acc2[k & 3], one, c1);
					if (MODE == 7) { // D I D I D
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						acc2[k & 3] = imad_add(acc2[k & 3], one, c1);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						acc2[(k + 1) & 3] = imad_add(acc2[(k + 1) & 3], one, c1);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
					}
					if (MODE == 8) { // D I D I D I
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						acc2[k & 3] = imad_add(acc2[k & 3], one, c1);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						acc2[(k + 1) & 3] = imad_add(acc2[(k + 1) & 3], one, c1);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
						acc2[(k + 2) & 3] = imad_add(acc2[(k + 2) & 3], one, c1);
					}
					if (MODE == 9) { // D D D + FADD
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
						facc[k & 3] = facc[k & 3] + fone;
					}
					if (MODE == 10) { // D D D + broadcast-pattern LDS (result summed by IADD3 every 4th)
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
						uint32_t t;
						asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(pb + ((rep * ILP + k) & 31) * 100));
						lacc = t;
					}
					if (MODE == 11) { // D D D + IMAD + broadcast LDS feeding the IMAD
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
						uint32_t t;
						asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(pb + ((rep * ILP + k) & 31) * 100));
						acc2[k & 3] = imad_add(acc2[k & 3], one, t);
					}
					if (MODE == 13 || MODE == 14 || MODE == 15) { // half2 max next to DPX
						if (MODE == 13 || MODE == 15)
							v[k] = __viaddmax_u16x2(v[k], c1, seed);
						if (MODE == 13) {
							x[k] = __viaddmax_u16x2(x[k], c1, seed);
							w[k] = __vimax3_u16x2(w[k], c1, seed);
						}
						uint32_t r;
						asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(hacc[k & 7]), "r"(c1));
						hacc[k & 7] = r;
					}
					if (MODE == 16 || MODE == 17) { // f32 3-input max next to DPX
						if (MODE == 16) {
							v[k] = __viaddmax_u16x2(v[k], c1, seed);
							x[k] = __viaddmax_u16x2(x[k], c1, seed);
							w[k] = __vimax3_u16x2(w[k], c1, seed);
						}
						facc[k & 3] = fmaxf(facc[k & 3], fone);
					}
					if (MODE == 12) { // D D D + plain add (ptxas picks IADD3 or IMAD.IADD)
						v[k] = __viaddmax_u16x2(v[k], c1, seed);
						x[k] = __viaddmax_u16x2(x[k], c1, seed);
						w[k] = __vimax3_u16x2(w[k], c1, seed);
						acc2[k & 3] += c1;
					}
					continue;
				}
				const uint32_t s = MODE >= 1 ? ps[((rep * ILP + k) & 63) * 32] : c1;
				const uint32_t y = __viaddmax_u16x2(v[k], c1, w[k]);
				x[k] = __viaddmax_u16x2(v[k], c1, x[k]);
				uint32_t d = MODE == 2 ? (v[k] ^ s) : imad_add(v[k], one, s);
				const uint32_t m = __vimax3_u16x2(d, x[k], y);
				v[k] = m;
				w[k] = y;
			}
		}
	}
	for (int k = 0; k < 8; k++)
		lacc ^= hacc[k];
	uint32_t acc = acc2[0] ^ acc2[1] ^ acc2[2] ^ acc2[3] ^ lacc ^ __float_as_uint(facc[0] + facc[1] + facc[2] + facc[3]);
	for (int k = 0; k < ILP; k++)
		acc ^= v[k] ^ w[k] ^ x[k];
	out[blockIdx.x * 128 + threadIdx.x] = acc;
}

template <int MODE, int ILP> static void run_mix(const char *name, uint32_t *out, int sms, double mhz, int occ_cap)
{
	auto kern = mixk<MODE, ILP>;
	cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
	int dyn = 0, occ = 0;
	cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, 0);
	while (occ_cap && occ > occ_cap) {
		dyn += 2048;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, dyn);
	}
	const int iters = 4096;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	for (int rep = 0; rep < 4; rep++) {
		cudaEventRecord(e0);
		kern<<<sms * occ, 128, dyn>>>(out, iters, 1u, 12345u + rep);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (rep && ms < best)
			best = ms;
	}
	const double cells_per_smsp = (double)occ * iters * 4.0 * ILP;
	printf("%-44s warps/SMSP %d  %.2f clk/cell/SMSP\n", name, occ, best * 1e-3 * mhz * 1e6 / cells_per_smsp);
}

// Library entry (bench.py): TCUPS ceiling of the bare register block for `method` (0 NW, 1 GA,
// 2 SW) and strip width (narrow = 16 columns / 5 blocks per SM, else 32 / 4) on `device`.
static void make_letters(std::vector<uint32_t> &h)
{
	uint32_t s = 12345;
	for (auto &w : h) {
		uint32_t v = 0;
		for (int b = 0; b < 4; b++) {
			s = s * 1664525u + 1013904223u;
			v |= (((s >> 16) % 20) * 4) << (8 * b);
		}
		w = v;
	}
}

extern "C" int sab_cell_peak(int method, int narrow, int device, double *tcups, double *clk_per_cell_pair)
{
	if (method < 0 || method > 2 || !tcups)
		return -1;
	if (cudaSetDevice(device) != cudaSuccess)
		return -2;
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device) != cudaSuccess)
		return -2;
	int khz = 0;
	cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device);
	const double mhz = khz / 1e3;
	uint32_t *out = nullptr, *letters = nullptr;
	std::vector<uint32_t> h(64 * 32);
	make_letters(h);
	if (cudaMalloc(&out, sizeof(uint32_t) * prop.multiProcessorCount * 16 * 128) != cudaSuccess ||
	    cudaMalloc(&letters, h.size() * 4) != cudaSuccess)
		return -3;
	cudaMemcpy(letters, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
	const int sms = prop.multiProcessorCount;
	double v = -1.0;
	const int key = method * 2 + (narrow ? 1 : 0);
	switch (key) {
	case 0: v = run<M_NW, 32, 4, true, true>(nullptr, out, letters, sms, mhz, 4, clk_per_cell_pair); break;
	case 1: v = run<M_NW, 16, 5, true, true>(nullptr, out, letters, sms, mhz, 5, clk_per_cell_pair); break;
	case 2: v = run<M_GA, 32, 4, true, true>(nullptr, out, letters, sms, mhz, 4, clk_per_cell_pair); break;
	case 3: v = run<M_GA, 16, 5, true, true>(nullptr, out, letters, sms, mhz, 5, clk_per_cell_pair); break;
	case 4: v = run<M_SW, 32, 4, true, true>(nullptr, out, letters, sms, mhz, 4, clk_per_cell_pair); break;
	default: v = run<M_SW, 16, 5, true, true>(nullptr, out, letters, sms, mhz, 5, clk_per_cell_pair); break;
	}
	cudaFree(out);
	cudaFree(letters);
	if (v < 0)
		return -4;
	*tcups = v;
	return 0;
}

#ifndef CELL_PEAK_LIB
int main()
{
	cudaDeviceProp prop;
	cudaGetDeviceProperties(&prop, 0);
	int khz = 0;
	cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
	const double mhz = khz / 1e3;
	uint32_t *out, *letters;
	cudaMalloc(&out, sizeof(uint32_t) * prop.multiProcessorCount * 16 * 128);
	std::vector<uint32_t> h(64 * 32);
	make_letters(h);
	cudaMalloc(&letters, h.size() * 4);
	cudaMemcpy(letters, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
	const int sms = prop.multiProcessorCount;
#define MIX(MODE, ILP, CAP) run_mix<MODE, ILP>("mix mode=" #MODE " (3 DDD, 13 DDD+HMNMX2, 14 HMNMX2, 15 D+HMNMX2, 16 DDD+FMNMX, 17 FMNMX) ilp=" #ILP, out, sms, mhz, CAP)
#ifdef WITH_MIX
	MIX(3, 8, 4);
	MIX(13, 8, 4);
	MIX(14, 8, 4);
	MIX(15, 8, 4);
	MIX(16, 8, 4);
	MIX(17, 8, 4);
#endif
#define RUN(M, K, B, L, R, NR, CAP) \
	run<M, K, B, L, R, NR>(#M " K=" #K " minb=" #B " lds=" #L " bar=" #R " NR=" #NR " cap=" #CAP, out, letters, sms, mhz, CAP)
	RUN(M_GA, 32, 4, true, false, 8, 0);
	RUN(M_SW, 32, 4, true, false, 8, 0);
	RUN(M_SW, 16, 5, true, false, 8, 0);
	RUN(M_SW, 16, 6, true, false, 8, 0);
	RUN(M_NW, 32, 4, true, false, 8, 0);
	RUN(M_NW, 32, 6, true, false, 8, 0);
	RUN(M_NW, 16, 8, true, false, 8, 0);
	return 0;
}
#endif // CELL_PEAK_LIB
