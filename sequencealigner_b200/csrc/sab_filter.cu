// sab_filter.cu -- redundancy filter decisions on the GPU (sab_filter).
//
// Reference semantics (src/bio/filter.c:14-62, evaluated sequentially, i.e. the
// deterministic `-T 1` result; the reference's OpenMP version races on lost[]):
//     for j = 1..N-1: lost[j] = exists i<j, !lost[i],
//                     (float)matches(i,j) / (float)min(len_i,len_j) >= threshold
// with matches = number of equal letters over the common prefix.
//
// B200 formulation: the O(N^2 L) letter comparisons are done brute force on the
// device, panel by panel over j; the order-dependent part (which i are still
// kept) is resolved on the host per panel:
//   * pairs with i BELOW the panel: kept[i] is final -> the kernel ORs a per-j flag;
//   * pairs with i INSIDE the panel: the kernel sets bit (j,i) of a panel bitmap,
//     the host walks j upwards and applies the recurrence.
// Letters are compared 16 at a time (uint4 loads of zero-padded, 16-byte aligned
// sequences, __vcmpeq4 + popc); the float32 divide/compare is IEEE (__fdiv_rn) so
// decisions are bit-identical to the host's.
#include "sab_internal.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <vector>

namespace {

constexpr int kFilterPanel = 2048; // j rows per pass; bitmap = 2048 x 2048 bits = 512 KB
constexpr int kJTile = 32;
constexpr int kIThreads = 128;

struct FilterParams {
	const uint4 *letters; // zero padded, sequence k starts at chunk off16[k]
	const int64_t *off16;
	const int32_t *len;
	const uint8_t *lost; // final for i < j0
	int j0, j1; // panel rows
	float thr;
	uint32_t *flag; // [j1 - j0] set when a kept i < j0 matches
	uint32_t *bits; // [(j1 - j0)][kFilterPanel / 32]
};

__global__ void __launch_bounds__(kIThreads) sab_filter_kernel(const FilterParams p)
{
	const int i = blockIdx.x * kIThreads + threadIdx.x;
	const int jt0 = p.j0 + blockIdx.y * kJTile;
	const int jt1 = min(jt0 + kJTile, p.j1);
	if (blockIdx.x * kIThreads >= jt1 - 1)
		return; // whole block above the diagonal
	const bool active = i < jt1 - 1 && !(i < p.j0 && p.lost[i]);
	const int li = i < p.j1 ? p.len[min(i, p.j1 - 1)] : 0;
	const uint4 *si = p.letters + (i < p.j1 ? p.off16[min(i, p.j1 - 1)] : 0);
	for (int j = max(jt0, 1); j < jt1; j++) {
		if (!active || i >= j)
			continue;
		const int lj = p.len[j];
		const uint4 *sj = p.letters + p.off16[j];
		const int ml = min(li, lj);
		const int nchunk = (ml + 15) >> 4;
		int m = 0;
		for (int c = 0; c < nchunk; c++) {
			const uint4 a = si[c], b = __ldg(sj + c);
			m += __popc(__vcmpeq4(a.x, b.x)) + __popc(__vcmpeq4(a.y, b.y)) +
			     __popc(__vcmpeq4(a.z, b.z)) + __popc(__vcmpeq4(a.w, b.w));
		}
		m >>= 3; // 8 bits per equal byte
		// zero padding that is equal in both sequences inside the last chunk
		m -= max(0, nchunk * 16 - max(li, lj));
		if (__fdiv_rn((float)m, (float)ml) >= p.thr) { // filter.c:52
			if (i < p.j0)
				p.flag[j - p.j0] = 1u;
			else
				atomicOr(&p.bits[(size_t)(j - p.j0) * (kFilterPanel / 32) +
						 ((i - p.j0) >> 5)],
					 1u << ((i - p.j0) & 31));
		}
	}
}

#define FCU(call)                                                                                \
	do {                                                                                     \
		cudaError_t e_ = (call);                                                         \
		if (e_ != cudaSuccess) {                                                         \
			rc = sab_internal_fail(ctx, SAB_ERR_CUDA, "%s: %s", #call,               \
					       cudaGetErrorString(e_));                          \
			goto done;                                                               \
		}                                                                                \
	} while (0)

} // namespace

extern "C" int sab_filter(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
			  float threshold, uint8_t *lost)
{
	if (!ctx || !seqs || !meta || !lost || num < 1)
		return SAB_ERR_ARG;
	memset(lost, 0, (size_t)num);
	if (threshold <= 0.0f) // filter.c:16
		return 0;
	const int dev = sab_internal_device(ctx);
	if (dev < 0)
		return sab_internal_fail(ctx, SAB_ERR_NODEVICE, "context has no device");

	std::vector<int64_t> off16(num);
	std::vector<int32_t> len(num);
	int64_t chunks = 0;
	for (int k = 0; k < num; k++) {
		if (meta[k].len < 1)
			return sab_internal_fail(ctx, SAB_ERR_ARG, "sequence %d is empty", k);
		off16[k] = chunks;
		len[k] = meta[k].len;
		chunks += (meta[k].len + 15) / 16;
	}
	std::vector<uint8_t> padded((size_t)chunks * 16, 0);
	for (int k = 0; k < num; k++)
		memcpy(padded.data() + off16[k] * 16, seqs + meta[k].off, (size_t)meta[k].len);

	int rc = SAB_OK, nlost = 0;
	uint4 *d_letters = nullptr;
	int64_t *d_off = nullptr;
	int32_t *d_len = nullptr;
	uint8_t *d_lost = nullptr;
	uint32_t *d_flag = nullptr, *d_bits = nullptr;
	const size_t bits_words = (size_t)kFilterPanel * (kFilterPanel / 32);
	std::vector<uint32_t> flag(kFilterPanel), bits(bits_words);

	FCU(cudaSetDevice(dev));
	FCU(cudaMalloc(&d_letters, padded.size()));
	FCU(cudaMalloc(&d_off, sizeof(int64_t) * num));
	FCU(cudaMalloc(&d_len, sizeof(int32_t) * num));
	FCU(cudaMalloc(&d_lost, (size_t)num));
	FCU(cudaMalloc(&d_flag, sizeof(uint32_t) * kFilterPanel));
	FCU(cudaMalloc(&d_bits, sizeof(uint32_t) * bits_words));
	FCU(cudaMemcpy(d_letters, padded.data(), padded.size(), cudaMemcpyHostToDevice));
	FCU(cudaMemcpy(d_off, off16.data(), sizeof(int64_t) * num, cudaMemcpyHostToDevice));
	FCU(cudaMemcpy(d_len, len.data(), sizeof(int32_t) * num, cudaMemcpyHostToDevice));
	FCU(cudaMemset(d_lost, 0, (size_t)num));

	for (int j0 = 0; j0 < num; j0 += kFilterPanel) {
		const int j1 = std::min(num, j0 + kFilterPanel);
		const int nj = j1 - j0;
		FCU(cudaMemset(d_flag, 0, sizeof(uint32_t) * kFilterPanel));
		FCU(cudaMemset(d_bits, 0, sizeof(uint32_t) * bits_words));
		FilterParams p{ d_letters, d_off, d_len, d_lost, j0, j1, threshold, d_flag, d_bits };
		dim3 grid((j1 - 1 + kIThreads - 1) / kIThreads, (nj + kJTile - 1) / kJTile);
		if (grid.x > 0) {
			sab_filter_kernel<<<grid, kIThreads>>>(p);
			FCU(cudaGetLastError());
		}
		FCU(cudaMemcpy(flag.data(), d_flag, sizeof(uint32_t) * nj, cudaMemcpyDeviceToHost));
		FCU(cudaMemcpy(bits.data(), d_bits, sizeof(uint32_t) * bits_words,
			       cudaMemcpyDeviceToHost));
		for (int j = j0; j < j1; j++) { // sequential resolve inside the panel
			bool l = flag[j - j0] != 0;
			const uint32_t *row = bits.data() + (size_t)(j - j0) * (kFilterPanel / 32);
			for (int w = 0; !l && w <= (j - j0) >> 5; w++) {
				uint32_t v = row[w];
				while (v) {
					const int i = j0 + w * 32 + __builtin_ctz(v);
					v &= v - 1;
					if (i < j && !lost[i]) {
						l = true;
						break;
					}
				}
			}
			if (j > 0 && l) {
				lost[j] = 1;
				nlost++;
			}
		}
		FCU(cudaMemcpy(d_lost + j0, lost + j0, (size_t)nj, cudaMemcpyHostToDevice));
	}
done:
	cudaFree(d_letters);
	cudaFree(d_off);
	cudaFree(d_len);
	cudaFree(d_lost);
	cudaFree(d_flag);
	cudaFree(d_bits);
	return rc ? rc : nlost;
}
