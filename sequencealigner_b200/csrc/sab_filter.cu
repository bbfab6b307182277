// sab_filter.cu -- redundancy filter decisions on the GPU (sab_filter, sab_filter_matches,
// sab_filter_resolve; sab_multi_filter in sab_capi.cu drives the same two phases on several devices).
//
// Reference semantics (src/bio/filter.c:14-62, evaluated sequentially, i.e. the
// deterministic `-T 1` result; the reference's OpenMP version races on lost[]):
//     for j = 1..N-1: lost[j] = exists i<j, !lost[i],
//                     (float)matches(i,j) / (float)min(len_i,len_j) >= threshold
// with matches = number of equal letters over the common prefix.
//
// B200 formulation, two phases:
//   A  (device, embarrassingly parallel, shards over devices by tiles of 32 j's)
//      every pair i<j whose prefix identity reaches the threshold is appended to a list --
//      regardless of lost[], so no tile depends on another.  Letters are compared 16 at a time
//      (uint4 loads of zero-padded, 16-byte aligned sequences); the float32 divide/compare of
//      filter.c:52 is folded into an integer table need[ml] = the smallest match count that
//      passes for a common prefix of ml letters, built on the host with the reference's own
//      expression, so a pair is dropped as soon as (matches so far + letters left) < need.
//   B  (host, sequential) the order-dependent recurrence over the listed pairs only.
// Near-duplicates are rare in real sets, so the list is short.  When it is not (a threshold most
// pairs pass) the list overflows and the exact panel-sequential path below takes over: panels of
// 2048 j's, kept[] final for everything below the panel, a bitmap for the pairs inside it.
#include "sab_internal.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <vector>

namespace {

constexpr int kFilterPanel = 2048; // j rows per pass of the sequential path; bitmap 2048 x 2048 bits
constexpr int kJTile = 32;
constexpr int kIThreads = 128;

double now_ms()
{
	using namespace std::chrono;
	return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// device buffers kept in the context between calls (grow-only)
struct FilterState {
	uint4 *letters = nullptr;
	int64_t *off16 = nullptr;
	int32_t *len = nullptr, *need = nullptr;
	uint8_t *lost = nullptr;
	uint32_t *flag = nullptr, *bits = nullptr;
	unsigned long long *list = nullptr, *count = nullptr;
	size_t cap_letters = 0, cap_num = 0, cap_need = 0, cap_list = 0;
};

void release_state(void *p)
{
	FilterState *st = static_cast<FilterState *>(p);
	cudaFree(st->letters);
	cudaFree(st->off16);
	cudaFree(st->len);
	cudaFree(st->need);
	cudaFree(st->lost);
	cudaFree(st->flag);
	cudaFree(st->bits);
	cudaFree(st->list);
	cudaFree(st->count);
	delete st;
}

struct FilterParams {
	const uint4 *letters; // zero padded, sequence k starts at chunk off16[k]
	const int64_t *off16;
	const int32_t *len;
	const int32_t *need; // [max_len + 1]
	int num;
	// phase A
	int shard, nshards;
	unsigned long long *list, *count;
	unsigned long long cap;
	// sequential path
	const uint8_t *lost; // final for i < j0
	int j0, j1;
	uint32_t *flag; // [j1 - j0] set when a kept i < j0 matches
	uint32_t *bits; // [(j1 - j0)][kFilterPanel / 32]
};

__device__ __forceinline__ int equal_bytes(const uint4 a, const uint4 b)
{
	return (__popc(__vcmpeq4(a.x, b.x)) + __popc(__vcmpeq4(a.y, b.y)) + __popc(__vcmpeq4(a.z, b.z)) +
		__popc(__vcmpeq4(a.w, b.w))) >> 3; // 8 bits per equal byte
}

// Does pair (i, j) reach the threshold?  si0 = first chunk of sequence i (in a register), sj0 that
// of j (from shared memory).  Stops as soon as the letters left cannot make up the difference.
__device__ __forceinline__ bool pair_passes(const FilterParams &p, const uint4 si0, const uint4 *si, int li,
					    const uint4 sj0, const uint4 *sj, int lj)
{
	const int ml = min(li, lj);
	const int need = __ldg(p.need + ml);
	const int nchunk = (ml + 15) >> 4;
	int m = equal_bytes(si0, sj0);
	for (int c = 1; c < nchunk; c++) {
		if (m + (ml - 16 * c) < need)
			return false; // out of reach: (float)(m + left) / (float)ml < threshold already
		m += equal_bytes(si[c], __ldg(sj + c));
	}
	// zero padding that is equal in both sequences inside the last chunk
	m -= max(0, nchunk * 16 - max(li, lj));
	return m >= need;
}

// Tile t of 32 j's belongs to shard (t % n), the direction flipping every round so the shards'
// loads (a tile costs ~ its j) stay level.
__device__ __forceinline__ int tile_of_shard(int y, int shard, int nshards)
{
	return y * nshards + ((y & 1) ? nshards - 1 - shard : shard);
}

__global__ void __launch_bounds__(kIThreads) sab_filter_pairs_kernel(const FilterParams p)
{
	__shared__ uint4 sj0[kJTile];
	__shared__ int slen[kJTile];
	const int tile = tile_of_shard(blockIdx.y, p.shard, p.nshards);
	const int jt0 = tile * kJTile;
	const int jt1 = min(jt0 + kJTile, p.num);
	if (jt0 >= p.num || (int)(blockIdx.x * kIThreads) >= jt1 - 1)
		return; // whole block on or above the diagonal
	if (threadIdx.x < kJTile) {
		const int j = min(jt0 + (int)threadIdx.x, p.num - 1);
		sj0[threadIdx.x] = p.letters[p.off16[j]];
		slen[threadIdx.x] = p.len[j];
	}
	__syncthreads();
	const int i = blockIdx.x * kIThreads + threadIdx.x;
	if (i >= jt1 - 1)
		return;
	const int li = p.len[i];
	const uint4 *si = p.letters + p.off16[i];
	const uint4 si0 = si[0];
	for (int j = max(jt0, i + 1); j < jt1; j++) {
		const uint4 *sj = p.letters + p.off16[j];
		if (pair_passes(p, si0, si, li, sj0[j - jt0], sj, slen[j - jt0])) {
			const unsigned long long k = atomicAdd(p.count, 1ull);
			if (k < p.cap)
				p.list[k] = ((unsigned long long)(unsigned)j << 32) | (unsigned)i;
		}
	}
}

__global__ void __launch_bounds__(kIThreads) sab_filter_panel_kernel(const FilterParams p)
{
	__shared__ uint4 sj0[kJTile];
	__shared__ int slen[kJTile];
	const int jt0 = p.j0 + blockIdx.y * kJTile;
	const int jt1 = min(jt0 + kJTile, p.j1);
	if ((int)(blockIdx.x * kIThreads) >= jt1 - 1)
		return;
	if (threadIdx.x < kJTile) {
		const int j = min(jt0 + (int)threadIdx.x, p.num - 1);
		sj0[threadIdx.x] = p.letters[p.off16[j]];
		slen[threadIdx.x] = p.len[j];
	}
	__syncthreads();
	const int i = blockIdx.x * kIThreads + threadIdx.x;
	if (i >= jt1 - 1 || (i < p.j0 && p.lost[i]))
		return;
	const int li = p.len[i];
	const uint4 *si = p.letters + p.off16[i];
	const uint4 si0 = si[0];
	for (int j = max(max(jt0, 1), i + 1); j < jt1; j++) {
		const uint4 *sj = p.letters + p.off16[j];
		if (!pair_passes(p, si0, si, li, sj0[j - jt0], sj, slen[j - jt0]))
			continue;
		if (i < p.j0)
			p.flag[j - p.j0] = 1u;
		else
			atomicOr(&p.bits[(size_t)(j - p.j0) * (kFilterPanel / 32) + ((i - p.j0) >> 5)],
				 1u << ((i - p.j0) & 31));
	}
}

#define FCU(call)                                                                                \
	do {                                                                                     \
		cudaError_t e_ = (call);                                                         \
		if (e_ != cudaSuccess)                                                           \
			return sab_internal_fail(ctx, e_ == cudaErrorMemoryAllocation ? SAB_ERR_NOMEM : SAB_ERR_CUDA, \
						 "%s: %s", #call, cudaGetErrorString(e_));       \
	} while (0)

template <typename T> int grow(sab_ctx *ctx, T **p, size_t *cap, size_t need)
{
	if (need <= *cap)
		return SAB_OK;
	cudaFree(*p);
	*p = nullptr;
	*cap = 0;
	FCU(cudaMalloc(p, need * sizeof(T)));
	*cap = need;
	return SAB_OK;
}

// need[ml] = the smallest number of matching letters that passes filter.c:52 for a common prefix
// of ml letters, found with the reference's own float32 expression ((float)m / (float)ml is
// monotone in m, so a bisection is exact).  need[ml] = ml + 1 when nothing passes.
void build_need(float threshold, int max_len, std::vector<int32_t> &need)
{
	need.assign((size_t)max_len + 1, 0);
	for (int ml = 1; ml <= max_len; ml++) {
		int lo = 0, hi = ml + 1; // first m in [0, ml] that passes, else ml + 1
		while (lo < hi) {
			const int mid = (lo + hi) / 2;
			const volatile float q = (float)mid / (float)ml;
			if (mid <= ml && q >= threshold)
				hi = mid;
			else
				lo = mid + 1;
		}
		need[ml] = lo;
	}
}

// Packs the letters (16-byte chunks, zero padded) and uploads them with the per-sequence tables.
int stage_sequences(sab_ctx *ctx, FilterState *st, const uint8_t *seqs, const sab_meta *meta, int32_t num,
		    float threshold, FilterParams *p)
{
	std::vector<int64_t> off16(num);
	std::vector<int32_t> len(num);
	int64_t chunks = 0;
	int max_len = 0;
	for (int k = 0; k < num; k++) {
		if (meta[k].len < 1 || meta[k].off < 0)
			return sab_internal_fail(ctx, SAB_ERR_ARG, "sequence %d has bad length/offset", k);
		off16[k] = chunks;
		len[k] = meta[k].len;
		max_len = std::max(max_len, meta[k].len);
		chunks += (meta[k].len + 15) / 16;
	}
	std::vector<uint8_t> padded((size_t)chunks * 16, 0);
	for (int k = 0; k < num; k++)
		memcpy(padded.data() + off16[k] * 16, seqs + meta[k].off, (size_t)meta[k].len);
	std::vector<int32_t> need;
	build_need(threshold, max_len, need);

	int rc;
	if ((rc = grow(ctx, &st->letters, &st->cap_letters, (size_t)chunks)))
		return rc;
	if (num > (int)st->cap_num) {
		size_t c1 = st->cap_num, c2 = st->cap_num, c3 = st->cap_num;
		if ((rc = grow(ctx, &st->off16, &c1, (size_t)num)) || (rc = grow(ctx, &st->len, &c2, (size_t)num)) ||
		    (rc = grow(ctx, &st->lost, &c3, (size_t)num)))
			return rc;
		st->cap_num = (size_t)num;
	}
	if ((rc = grow(ctx, &st->need, &st->cap_need, need.size())))
		return rc;
	if (!st->count)
		FCU(cudaMalloc(&st->count, sizeof(unsigned long long)));
	FCU(cudaMemcpy(st->letters, padded.data(), padded.size(), cudaMemcpyHostToDevice));
	FCU(cudaMemcpy(st->off16, off16.data(), sizeof(int64_t) * num, cudaMemcpyHostToDevice));
	FCU(cudaMemcpy(st->len, len.data(), sizeof(int32_t) * num, cudaMemcpyHostToDevice));
	FCU(cudaMemcpy(st->need, need.data(), sizeof(int32_t) * need.size(), cudaMemcpyHostToDevice));
	*p = FilterParams{};
	p->letters = st->letters;
	p->off16 = st->off16;
	p->len = st->len;
	p->need = st->need;
	p->num = num;
	return SAB_OK;
}

FilterState *state_of(sab_ctx *ctx)
{
	void **slot = sab_internal_filter_slot(ctx, release_state);
	if (!*slot)
		*slot = new FilterState;
	return static_cast<FilterState *>(*slot);
}

// The exact panel-sequential path (any density of matches): sequences already staged.
int filter_sequential(sab_ctx *ctx, FilterState *st, FilterParams p, int32_t num, uint8_t *lost)
{
	const size_t bits_words = (size_t)kFilterPanel * (kFilterPanel / 32);
	if (!st->flag)
		FCU(cudaMalloc(&st->flag, sizeof(uint32_t) * kFilterPanel));
	if (!st->bits)
		FCU(cudaMalloc(&st->bits, sizeof(uint32_t) * bits_words));
	std::vector<uint32_t> flag(kFilterPanel), bits(bits_words);
	FCU(cudaMemset(st->lost, 0, (size_t)num));
	int nlost = 0;
	p.lost = st->lost;
	p.flag = st->flag;
	p.bits = st->bits;
	for (int j0 = 0; j0 < num; j0 += kFilterPanel) {
		const int j1 = std::min(num, j0 + kFilterPanel);
		const int nj = j1 - j0;
		FCU(cudaMemset(st->flag, 0, sizeof(uint32_t) * kFilterPanel));
		FCU(cudaMemset(st->bits, 0, sizeof(uint32_t) * bits_words));
		p.j0 = j0;
		p.j1 = j1;
		dim3 grid((j1 - 1 + kIThreads - 1) / kIThreads, (nj + kJTile - 1) / kJTile);
		if (grid.x > 0) {
			sab_filter_panel_kernel<<<grid, kIThreads>>>(p);
			FCU(cudaGetLastError());
		}
		FCU(cudaMemcpy(flag.data(), st->flag, sizeof(uint32_t) * nj, cudaMemcpyDeviceToHost));
		FCU(cudaMemcpy(bits.data(), st->bits, sizeof(uint32_t) * bits_words, cudaMemcpyDeviceToHost));
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
		FCU(cudaMemcpy(st->lost + j0, lost + j0, (size_t)nj, cudaMemcpyHostToDevice));
	}
	return nlost;
}

} // namespace

int sab_internal_filter_matches(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
				float threshold, int shard, int nshards, std::vector<uint64_t> &pairs,
				int *overflow)
{
	pairs.clear();
	*overflow = 0;
	const int dev = sab_internal_device(ctx);
	if (dev < 0)
		return sab_internal_fail(ctx, SAB_ERR_NODEVICE, "context has no device");
	if (nshards < 1 || shard < 0 || shard >= nshards)
		return sab_internal_fail(ctx, SAB_ERR_ARG, "bad shard %d of %d", shard, nshards);
	if (threshold <= 0.0f || num < 2) // filter.c:16
		return SAB_OK;
	FCU(cudaSetDevice(dev));
	FilterState *st = state_of(ctx);
	FilterParams p;
	int rc = stage_sequences(ctx, st, seqs, meta, num, threshold, &p);
	if (rc)
		return rc;
	int flags = 0;
	sab_internal_flags(ctx, &flags);
	const size_t cap = (flags & SAB_FLAG_TEST_FILTER_CAP) ?
				   16 :
				   std::min<size_t>(32u << 20, std::max<size_t>(1u << 20, (size_t)num * 64));
	if ((rc = grow(ctx, &st->list, &st->cap_list, cap)))
		return rc;
	FCU(cudaMemset(st->count, 0, sizeof(unsigned long long)));
	p.shard = shard;
	p.nshards = nshards;
	p.list = st->list;
	p.count = st->count;
	p.cap = cap;
	const int tiles = (num + kJTile - 1) / kJTile;
	const int my_tiles = (tiles - 1) / nshards + 1; // rounds; the kernel skips tiles past the end
	dim3 grid((num - 1 + kIThreads - 1) / kIThreads, my_tiles);
	sab_filter_pairs_kernel<<<grid, kIThreads>>>(p);
	FCU(cudaGetLastError());
	unsigned long long count = 0;
	FCU(cudaMemcpy(&count, st->count, sizeof count, cudaMemcpyDeviceToHost));
	if (count > cap) {
		*overflow = 1;
		return SAB_OK;
	}
	pairs.resize((size_t)count);
	if (count)
		FCU(cudaMemcpy(pairs.data(), st->list, sizeof(uint64_t) * count, cudaMemcpyDeviceToHost));
	return SAB_OK;
}

extern "C" {

int sab_filter_matches(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num, float threshold,
		       int32_t shard, int32_t nshards, uint64_t *pairs, int64_t cap, int64_t *count)
{
	if (!ctx || !seqs || !meta || !count || num < 1 || (cap > 0 && !pairs))
		return SAB_ERR_ARG;
	std::vector<uint64_t> v;
	int overflow = 0;
	const double t0 = now_ms();
	const int rc = sab_internal_filter_matches(ctx, seqs, meta, num, threshold, shard, nshards, v, &overflow);
	if (rc)
		return rc;
	*count = (int64_t)v.size();
	if (overflow || (int64_t)v.size() > cap) {
		*count = overflow ? -1 : (int64_t)v.size();
		return sab_internal_fail(ctx, SAB_ERR_OVERFLOW,
					 "more matching pairs than the list holds; use sab_filter (exact "
					 "panel-sequential path)");
	}
	if (!v.empty())
		memcpy(pairs, v.data(), v.size() * sizeof(uint64_t));
	sab_internal_filter_stats(ctx, now_ms() - t0, (int64_t)v.size());
	return SAB_OK;
}

int sab_filter_resolve(int32_t num, const uint64_t *pairs, int64_t count, uint8_t *lost)
{
	if (num < 1 || !lost || count < 0 || (count > 0 && !pairs))
		return SAB_ERR_ARG;
	memset(lost, 0, (size_t)num);
	// bucket the listed pairs by j (counting sort), then walk j upwards: filter.c:35-55 restricted
	// to the pairs that pass the threshold
	std::vector<int64_t> start((size_t)num + 1, 0);
	for (int64_t k = 0; k < count; k++) {
		const uint32_t j = (uint32_t)(pairs[k] >> 32), i = (uint32_t)pairs[k];
		if (j >= (uint32_t)num || i >= j)
			return SAB_ERR_ARG;
		start[j + 1]++;
	}
	for (int j = 0; j < num; j++)
		start[j + 1] += start[j];
	std::vector<uint32_t> by_j((size_t)count);
	std::vector<int64_t> fill(start.begin(), start.end() - 1);
	for (int64_t k = 0; k < count; k++)
		by_j[fill[pairs[k] >> 32]++] = (uint32_t)pairs[k];
	int nlost = 0;
	for (int j = 1; j < num; j++)
		for (int64_t k = start[j]; k < start[j + 1]; k++)
			if (!lost[by_j[k]]) {
				lost[j] = 1;
				nlost++;
				break;
			}
	return nlost;
}

int sab_filter(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num, float threshold,
	       uint8_t *lost)
{
	if (!ctx || !seqs || !meta || !lost || num < 1)
		return SAB_ERR_ARG;
	memset(lost, 0, (size_t)num);
	if (threshold <= 0.0f) // filter.c:16
		return 0;
	const double t0 = now_ms();
	int flags = 0;
	sab_internal_flags(ctx, &flags);
	std::vector<uint64_t> v;
	int overflow = 0;
	if (!(flags & SAB_FLAG_FILTER_SEQUENTIAL)) {
		const int rc = sab_internal_filter_matches(ctx, seqs, meta, num, threshold, 0, 1, v, &overflow);
		if (rc)
			return rc;
		if (!overflow) {
			const int nlost = sab_filter_resolve(num, v.data(), (int64_t)v.size(), lost);
			sab_internal_filter_stats(ctx, now_ms() - t0, (int64_t)v.size());
			return nlost;
		}
	}
	const int dev = sab_internal_device(ctx);
	if (dev < 0)
		return sab_internal_fail(ctx, SAB_ERR_NODEVICE, "context has no device");
	FCU(cudaSetDevice(dev));
	FilterState *st = state_of(ctx);
	FilterParams p;
	int rc = stage_sequences(ctx, st, seqs, meta, num, threshold, &p);
	if (rc)
		return rc;
	const int nlost = filter_sequential(ctx, st, p, num, lost);
	sab_internal_filter_stats(ctx, now_ms() - t0, -1);
	return nlost;
}

} // extern "C"
