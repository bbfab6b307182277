// sab_noalloc.cu -- sab_sw_pairs_noalloc: Smith-Waterman scores of a caller-given list of pairs WITHOUT any
// allocation (include/seqalign_b200.h).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it).
//
// The kernel is a variant of sab_rerun_sw_kernel (sab_kernels.cuh): one thread per pair, int32 arithmetic
// exactly as src/bio/method/sw.c:15-61, two rolling DP rows (M and Y) per thread in global scratch,
// interleaved across threads (element c of thread t at [c * threads + t]: a warp's accesses coalesce).  The
// differences: it reads the caller's RAW sequence store (device copies of struct input's `seqs` / `meta`,
// src/io/input.h:6-11) instead of the packed layouts of sab_upload, and its scratch is not cudaMalloc'ed per
// context but lives in ONE file-scope array of fixed size:
//
//   g_noalloc_scratch   16 MiB, static __device__: every GPU that loads the library holds it until exit.
//
// In host code that name is a host-side stand-in, not a device address: the address the kernel gets comes
// from cudaGetSymbolAddress, looked up once per device with that device current (scratch_of_current_device).
// All calls on a device share the array and it keeps the last call's contents.  Nothing in it has to start at
// zero -- a thread writes the border row of a pair before it reads it -- so nothing is cleared; but two calls
// that OVERLAP on one GPU (different streams, or a captured graph replayed concurrently with itself) race on
// it: the caller serialises them.
#include "../../include/seqalign_b200.h"

#include <cuda_runtime.h>

#include <climits>
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace sab {

constexpr int kNoallocBlock = 128;
constexpr int kNoallocMaxDevices = 64;
constexpr size_t kNoallocWords = SAB_NOALLOC_SCRATCH_BYTES / sizeof(int32_t);

static __device__ __align__(256) int32_t g_noalloc_scratch[kNoallocWords];

struct NoallocParams {
	const uint8_t *seqs; // device: NUL-separated upper-case letters
	const sab_meta *meta; // device: [num] (offset, length)
	const int32_t *pair_i, *pair_j; // device: [count] original indices, pair_i[k] != pair_j[k]
	int32_t *scores; // device: [count]
	int64_t count;
	int32_t num, max_len;
	int32_t gap_open, gap_ext; // <= 0
	int32_t *scratch; // == address of g_noalloc_scratch on this device
	int8_t lut[128]; // SEQ_LUT: ASCII -> 0..23, -1 = not a letter
	int16_t sub[24 * 24]; // SUB_MAT
};

// One pair: rows = letters of sequence `a`, columns = letters of sequence `b` (the tables are symmetric and
// so is the score).  m / y = this thread's rolling rows, `st` words apart per column.  Returns the running
// maximum of M (sw.c:32,53-57), or INT32_MIN when a letter is outside the alphabet.
__host__ __device__ inline int32_t noalloc_pair_sw(const uint8_t *a, const int la, const uint8_t *b, const int lb,
						   int32_t *m, int32_t *y, const int64_t st, const int o, const int e,
						   const int8_t *lut, const int16_t *sub)
{
	const int NEG = INT_MIN / 2; // SCORE_MIN, src/bio/align.h:19
	for (int c = 0; c <= lb; c++) { // sw.c:18-24
		m[c * st] = 0;
		y[c * st] = NEG;
	}
	int best = 0;
	for (int r = 0; r < la; r++) {
		const int ar = a[r] < 128 ? lut[a[r]] : -1;
		if (ar < 0)
			return INT_MIN;
		const int16_t *srow = sub + ar * 24;
		int diag = m[0], left = 0, x = NEG; // column 0: M = 0, X = -inf (sw.c:26-30)
		for (int c = 1; c <= lb; c++) {
			const int bc = b[c - 1] < 128 ? lut[b[c - 1]] : -1;
			if (bc < 0)
				return INT_MIN;
			const int up = m[c * st];
			const int xo = left + o, xe = x + e;
			x = xo > xe ? xo : xe; // sw.c:39-44
			const int yo = up + o, ye = y[c * st] + e;
			const int yy = yo > ye ? yo : ye; // sw.c:45-50
			int v = diag + srow[bc]; // sw.c:51-57
			v = v > 0 ? v : 0;
			v = v > x ? v : x;
			v = v > yy ? v : yy;
			diag = up;
			m[c * st] = v;
			y[c * st] = yy;
			left = v;
			best = best > v ? best : v;
		}
	}
	return best;
}

__global__ void __launch_bounds__(kNoallocBlock) sab_sw_pairs_static_kernel(const NoallocParams p)
{
	__shared__ int8_t s_lut[128];
	__shared__ int16_t s_sub[24 * 24];
	for (int i = threadIdx.x; i < 128; i += blockDim.x)
		s_lut[i] = p.lut[i];
	for (int i = threadIdx.x; i < 24 * 24; i += blockDim.x)
		s_sub[i] = p.sub[i];
	__syncthreads();
	const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
	const int64_t me = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	int32_t *m = p.scratch + me; // [2][max_len + 1][nthreads]
	int32_t *y = m + (int64_t)(p.max_len + 1) * nthreads;
	for (int64_t k = me; k < p.count; k += nthreads) {
		const int i = p.pair_i[k], j = p.pair_j[k];
		int32_t score = INT_MIN; // bad index / length beyond max_len / letter outside the alphabet
		if (i >= 0 && j >= 0 && i < p.num && j < p.num && i != j) {
			const sab_meta mi = p.meta[i], mj = p.meta[j];
			if (mi.len >= 1 && mj.len >= 1 && mi.len <= p.max_len && mj.len <= p.max_len && mi.off >= 0 &&
			    mj.off >= 0)
				score = noalloc_pair_sw(p.seqs + mj.off, mj.len, p.seqs + mi.off, mi.len, m, y, nthreads,
							p.gap_open, p.gap_ext, s_lut, s_sub);
		}
		p.scores[k] = score;
	}
}

namespace {

int noalloc_fail(char *err, size_t cap, int code, const char *fmt, ...)
{
	if (err && cap) {
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(err, cap, fmt, ap);
		va_end(ap);
	}
	return code;
}

// Device address of g_noalloc_scratch on the CURRENT device: looked up once per device, with that device
// current (a symbol has one address per device; the host-side name is not one of them).
cudaError_t scratch_of_current_device(int32_t **out, int *device)
{
	static std::mutex mu;
	static int32_t *ptr[kNoallocMaxDevices] = { nullptr };
	cudaError_t e = cudaGetDevice(device);
	if (e != cudaSuccess)
		return e;
	if (*device < 0 || *device >= kNoallocMaxDevices)
		return cudaErrorInvalidDevice;
	std::lock_guard<std::mutex> lock(mu);
	if (!ptr[*device]) {
		void *p = nullptr;
		e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess)
			return e;
		ptr[*device] = static_cast<int32_t *>(p);
	}
	*out = ptr[*device];
	return cudaSuccess;
}

bool on_device(const void *p)
{
	cudaPointerAttributes at{};
	if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
		cudaGetLastError();
		return false;
	}
	return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

} // namespace
} // namespace sab

extern "C" int sab_sw_pairs_noalloc(const sab_config *cfg, const uint8_t *d_seqs, const sab_meta *d_meta, int32_t num,
				    int32_t max_len, const int32_t *d_pair_i, const int32_t *d_pair_j, int64_t count,
				    int32_t *d_scores, void *stream, char *err, size_t err_cap)
{
	using namespace sab;
	if (!cfg)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "null configuration");
	if (cfg->method != SAB_SW)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "sab_sw_pairs_noalloc scores Smith-Waterman only (method %d given)",
				    cfg->method);
	if (cfg->gap_open > 0 || cfg->gap_ext > 0)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "gap values must be stored <= 0 (align.c:127)");
	if (count < 0 || num < 0 || max_len < 1)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "count, num must be >= 0 and max_len >= 1");
	if (max_len > SAB_NOALLOC_MAX_LEN)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG,
				    "max_len %d is beyond SAB_NOALLOC_MAX_LEN = %d: the static scratch (%u bytes) holds two "
				    "rows of at most that many columns for %d threads; use sab_align for longer sequences",
				    max_len, SAB_NOALLOC_MAX_LEN, (unsigned)SAB_NOALLOC_SCRATCH_BYTES,
				    (int)(kNoallocWords / (2 * (size_t)(SAB_NOALLOC_MAX_LEN + 1))));
	NoallocParams p{};
	for (int c = 0; c < 128; c++) {
		const int32_t v = cfg->seq_lut[c];
		if (v >= 24)
			return noalloc_fail(err, err_cap, SAB_ERR_ARG, "seq_lut[%d] = %d is not a letter index", c, v);
		p.lut[c] = (int8_t)(v < 0 ? -1 : v);
	}
	for (int a = 0; a < 24; a++)
		for (int b = 0; b < 24; b++) {
			const int32_t v = cfg->sub_mat[a * 24 + b];
			if (v != cfg->sub_mat[b * 24 + a])
				return noalloc_fail(err, err_cap, SAB_ERR_ARG, "substitution matrix is not symmetric");
			if (v < -32768 || v > 32767)
				return noalloc_fail(err, err_cap, SAB_ERR_ARG, "substitution score %d out of the int16 range", v);
			p.sub[a * 24 + b] = (int16_t)v;
		}
	// the int32 range of the reference (sab_create applies the same guard through int32_safe)
	int smax = 0;
	for (int k = 0; k < 24 * 24; k++)
		smax = cfg->sub_mat[k] > smax ? cfg->sub_mat[k] : smax;
	if ((int64_t)smax * max_len >= (1ll << 29) || -(int64_t)cfg->gap_open >= (1ll << 28) ||
	    -(int64_t)cfg->gap_ext >= (1ll << 28))
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "lengths x scores exceed the int32 score range");

	// set-up, once per device (a call with count == 0 does just this: make it before a graph capture)
	int device = -1;
	int32_t *scratch = nullptr;
	cudaError_t e = scratch_of_current_device(&scratch, &device);
	if (e != cudaSuccess) {
		cudaGetLastError();
		return noalloc_fail(err, err_cap, e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? SAB_ERR_NODEVICE : SAB_ERR_CUDA,
				    "cudaGetSymbolAddress(g_noalloc_scratch): %s", cudaGetErrorString(e));
	}
	if (count == 0)
		return SAB_OK;
	if (!d_seqs || !d_meta || !d_pair_i || !d_pair_j || !d_scores)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "null buffer");
	if (!on_device(d_seqs) || !on_device(d_meta) || !on_device(d_pair_i) || !on_device(d_pair_j) || !on_device(d_scores))
		return noalloc_fail(err, err_cap, SAB_ERR_ARG,
				    "seqs, meta, pair lists and scores must be DEVICE buffers of the current device (%d): this "
				    "entry point copies and allocates nothing", device);

	// as many threads as the scratch holds rows for, and no more than there are pairs
	const int64_t per_thread = 2 * (int64_t)(max_len + 1);
	int64_t threads = (int64_t)kNoallocWords / per_thread / kNoallocBlock * kNoallocBlock;
	if (threads < kNoallocBlock)
		return noalloc_fail(err, err_cap, SAB_ERR_ARG, "max_len %d leaves no room for one block in the static scratch", max_len);
	const int64_t wanted = (count + kNoallocBlock - 1) / kNoallocBlock * kNoallocBlock;
	threads = threads < wanted ? threads : wanted;
	threads = threads < (1 << 16) ? threads : (1 << 16);
	p.seqs = d_seqs;
	p.meta = d_meta;
	p.pair_i = d_pair_i;
	p.pair_j = d_pair_j;
	p.scores = d_scores;
	p.count = count;
	p.num = num;
	p.max_len = max_len;
	p.gap_open = cfg->gap_open;
	p.gap_ext = cfg->gap_ext;
	p.scratch = scratch;
	sab_sw_pairs_static_kernel<<<(unsigned)(threads / kNoallocBlock), kNoallocBlock, 0, static_cast<cudaStream_t>(stream)>>>(p);
	e = cudaGetLastError();
	if (e != cudaSuccess)
		return noalloc_fail(err, err_cap, SAB_ERR_CUDA, "sab_sw_pairs_static_kernel launch: %s", cudaGetErrorString(e));
	return SAB_OK; // asynchronous on `stream`, like a kernel launch
}
