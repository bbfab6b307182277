// sab_capi.cu -- C ABI of libseqalign_b200.so (include/seqalign_b200.h).
//
// Host side of the B200 score-only all-vs-all path: context, upload packing
// (length-binned SoA), row-strip scheduling over the triangular pair space,
// launch of the sm_100a kernels (sab_kernels.cuh) and tiled write-back into the
// caller's similarity matrix.  Replaces the reference's device backend
// (reference: src/interface/seqalign_cuda.c:95-296) -- one launch per row strip
// instead of 64 Mi-pair linear batches, nothing N^2-sized is required on the
// device beyond two strip buffers.
//
// There is no CPU fallback in this file: without a usable sm_100 device every
// compute entry point returns SAB_ERR_NODEVICE / SAB_ERR_CUDA.
#define SAB_WITH_RERUN_KERNEL
#include "sab_module.h"
#include "sab_wave.cuh" // WaveSmem<K> (shared-memory size per width); the kernels themselves are device images

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

namespace {

using sab::DeviceSeqs;
using sab::TileParams;

constexpr int kPanelQuantum = 128; // explicit panel heights are multiples of 4 query groups
constexpr double kTailBelowCellsPerPair = 12000.0; // panel_bounds: geometric tail below this mean len^2
constexpr size_t kStageBytes = 32u << 20; // pinned bounce buffer (x2)
constexpr unsigned kRerunCap = 1u << 20;
constexpr int kRerunBlock = 128;

double now_ms()
{
	using namespace std::chrono;
	return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

inline int64_t tri(int64_t j)
{
	return j * (j - 1) / 2; // == alignments(j), reference: src/util/macros.h:13
}

struct StripBuf {
	int32_t *blk = nullptr;
	int32_t *blkT = nullptr;
	size_t blk_cap = 0, blkT_cap = 0; // elements
	uint2 *bnd = nullptr;
	size_t bnd_cap = 0;
	unsigned *counter = nullptr; // [0] work counter, [1] rerun count
	int *chunk_ids = nullptr; // per-width lists of subject chunks of the strip in flight
	size_t chunk_ids_cap = 0;
	std::vector<int> chunk_host;
	unsigned long long *rerun_list = nullptr;
	int32_t *rerun_scratch = nullptr;
	size_t rerun_scratch_cap = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev_start = nullptr, ev_done = nullptr;
	// the launches of one strip (one per register-strip width in use) are independent: they run on
	// side streams forked from `stream`, so the tail of one overlaps the start of the next
	cudaStream_t side[15] = { nullptr };
	cudaEvent_t ev_fork = nullptr, ev_join[15] = { nullptr };
};

constexpr int kCounterWords = 32; // [1] re-run count, [4 + w] work counter of the width-w launch

} // namespace

struct sab_ctx {
	sab_config cfg{};
	std::string err;
	int sm_count = 0;
	// scoring
	int32_t sub25[sab::kLetters * sab::kLetters]{};
	int min_s = 0, max_s = 0;
	int gapA = 0, gapE = 0;
	int32_t *d_sub25 = nullptr;
	// sequence store (host copies of the small tables)
	bool uploaded = false;
	int32_t num = 0, max_len = 0;
	std::vector<int32_t> panel_start; // [npanels + 1] row boundaries of the row strips (panel_bounds)
	std::vector<int32_t> slot_len, slot_orig;
	std::vector<int64_t> len_prefix; // len_prefix[s] = sum of slot_len[0..s)
	std::vector<int64_t> pad16_prefix, pad32_prefix; // same, lengths rounded up to 16 / 32 columns
	std::vector<int32_t> panel_maxlen;
	int32_t *d_slot_len = nullptr, *d_slot_orig = nullptr;
	int64_t *d_col_off = nullptr, *d_grp_off = nullptr;
	uint32_t *d_col_words = nullptr;
	int col_bits = 5;
	uint32_t col_dec = 0;
	uint32_t *d_rowpack = nullptr;
	size_t cap_seqs[6]{}; // bytes allocated behind the six pointers above (reused across uploads)
	// execution
	StripBuf buf[2];
	void *stage[2] = { nullptr, nullptr };
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t ev_copy[2] = { nullptr, nullptr };
	cudaEvent_t ev_span[2] = { nullptr, nullptr }; // first kernel start / last kernel end
	struct Variant { // [method][packed][width index], set up on first use (wave_variant)
		void *fn = nullptr; // CUfunction of the specialised image (sab_module.cu)
		int smem = 0, grid = 0;
	} variants[3][2][16];
	void *modules = nullptr; // sab_module.cu cache of this context's loaded images
	bool sw_prefer_int32 = false; // a strip's re-run list overflowed: later risky strips start in int32
	sab_stats stats{};
	// sab_filter.cu keeps its device buffers here (released by sab_destroy)
	void *filter_state = nullptr;
	void (*filter_free)(void *) = nullptr;
};

namespace {

int fail(sab_ctx *c, int code, const char *fmt, ...)
{
	char b[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(b, sizeof b, fmt, ap);
	va_end(ap);
	if (c)
		c->err = b;
	return code;
}

#define CU(call)                                                                              \
	do {                                                                                  \
		cudaError_t e_ = (call);                                                      \
		if (e_ != cudaSuccess)                                                        \
			return fail(ctx, e_ == cudaErrorMemoryAllocation ? SAB_ERR_NOMEM :    \
									   SAB_ERR_CUDA,      \
				    "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__,  \
				    __LINE__);                                                \
	} while (0)

// grow-only device buffer
int ensure(sab_ctx *ctx, void **p, size_t *cap, size_t need, size_t elem)
{
	if (need <= *cap)
		return SAB_OK;
	cudaFree(*p);
	*p = nullptr;
	*cap = 0;
	cudaError_t e = cudaMalloc(p, need * elem);
	if (e != cudaSuccess)
		return fail(ctx, SAB_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", need * elem,
			    cudaGetErrorString(e));
	*cap = need;
	return SAB_OK;
}

// Register-strip widths (DP columns a warp holds in registers).  The kernel exists once per width
// (one device image each, sab_wave_inst.cu / sab_module.cu); a row strip runs at the width that pads
// its subjects least (choose_widths) instead of a global 32 or 16.  Widest first; int32 strips use
// the first and the last entry only.
struct WidthInfo {
	int k;
	int smem; // sizeof(WaveSmem<K>)
};
#define SAB_WIDTH_ENTRY(K) { K, (int)sizeof(sab::WaveSmem<K>) },
constexpr WidthInfo kWidths[] = { SAB_WIDTHS(SAB_WIDTH_ENTRY) };
#undef SAB_WIDTH_ENTRY
constexpr int kNumWidths = (int)(sizeof(kWidths) / sizeof(kWidths[0]));
constexpr int kWidest = 0, kNarrowest = kNumWidths - 1;
static_assert(kNumWidths <= 15 && kWidths[kWidest].k == sab::kStrip, "width table");

// Relative cost of one DP column in a strip of width K: a step carries the same bookkeeping
// (~165 instructions) whatever its width.  Calibrated on B200: a 16-column step costs 5.5 % more
// per column than a 32-column one (profiles/r1_cell_shape_ceiling.md) -> 1 + 1.87 / K.
inline double column_cost(int k)
{
	return 1.0 + 1.87 / k;
}

typedef sab_ctx::Variant WaveVariant;

// The per-lane gap constant of the add-max ops (sab_wave.cuh: CellK::lG), written into the image
// as an immediate operand: GA o - e, SW e; NW has none.
uint32_t lane_gap_constant(const sab_ctx *ctx, bool packed)
{
	const int v = ctx->cfg.method == SAB_GA ? ctx->gapA - ctx->gapE : ctx->gapE;
	return packed ? ((uint32_t)v & 0xffffu) * 0x10001u : (uint32_t)v;
}

// One kernel variant, set up on first use: its device image is specialised for this context's gap
// constant and loaded (sab_module.cu, ~1 ms; a job uses a handful of the 24).
int wave_variant(sab_ctx *ctx, int method, bool packed, int widx, WaveVariant *out)
{
	WaveVariant &v = ctx->variants[method][packed][widx];
	if (!v.fn) {
		const int k = kWidths[widx].k;
		if (!packed && widx != kWidest && widx != kNarrowest)
			return fail(ctx, SAB_ERR_STATE, "no int32 kernel of width %d", k);
		int occ = 0;
		void *fn = nullptr;
		std::string err;
		// cbnd, the last member of WaveSmem, belongs to the hoisted boundary exchange (sab_wave.cuh):
		// methods built without it (SW) do not get the 2 KB
		const int smem = kWidths[widx].smem -
				 (SAB_HOIST_METHODS(method) ? 0 : (int)sizeof(sab::WaveSmem<sab::kStrip>::cbnd));
		const int rc = sab_module_get(&ctx->modules, method, packed, k, lane_gap_constant(ctx, packed),
					      smem, &fn, &occ, &err);
		if (rc)
			return fail(ctx, rc, "%s", err.c_str());
		if (occ < 1)
			return fail(ctx, SAB_ERR_CUDA, "kernel does not fit on an SM");
		v.smem = smem;
		v.grid = occ * ctx->sm_count;
		v.fn = fn;
	}
	*out = v;
	return SAB_OK;
}

void free_seqs(sab_ctx *c)
{
	cudaFree(c->d_slot_len);
	cudaFree(c->d_slot_orig);
	cudaFree(c->d_col_off);
	cudaFree(c->d_grp_off);
	cudaFree(c->d_col_words);
	cudaFree(c->d_rowpack);
	c->d_slot_len = c->d_slot_orig = nullptr;
	c->d_col_off = c->d_grp_off = nullptr;
	c->d_col_words = nullptr;
	c->d_rowpack = nullptr;
	for (size_t &cap : c->cap_seqs)
		cap = 0;
	c->uploaded = false;
}

void free_bufs(sab_ctx *c)
{
	for (auto &b : c->buf) {
		cudaFree(b.blk);
		cudaFree(b.blkT);
		cudaFree(b.bnd);
		cudaFree(b.rerun_scratch);
		cudaFree(b.chunk_ids);
		b.chunk_ids = nullptr;
		b.chunk_ids_cap = 0;
		b.blk = b.blkT = nullptr;
		b.bnd = nullptr;
		b.rerun_scratch = nullptr;
		b.blk_cap = b.blkT_cap = b.bnd_cap = b.rerun_scratch_cap = 0;
	}
}

// Parallel host copy of one staged chunk into the (pageable, possibly file-backed)
// destination matrix: plain stores only (SURVEY section 8b "Output").
struct Seg {
	int32_t *dst;
	const int32_t *src;
	size_t n;
};

// Host threads a context may use for packing and for the copy into the caller's matrix:
// sab_config.host_threads (the reference's -T), else min(cores, 8).
int host_thread_cap(const sab_ctx *ctx)
{
	if (ctx && ctx->cfg.host_threads > 0)
		return ctx->cfg.host_threads;
	return (int)std::max(1u, std::min(std::thread::hardware_concurrency(), 8u));
}

// Runs fn(first, last) over [0, n) split into contiguous ranges on up to `cap` host threads
// (`work` = total units, to stay single-threaded for small inputs).
template <typename F> void parallel_ranges(int cap, int32_t n, int64_t work, F fn)
{
	const int nthreads = (int)std::min<int64_t>(cap, work / (1 << 19) + 1);
	if (nthreads <= 1 || n < nthreads) {
		fn(0, n);
		return;
	}
	std::vector<std::thread> th;
	for (int t = 0; t < nthreads; t++)
		th.emplace_back([=]() { fn((int32_t)((int64_t)n * t / nthreads), (int32_t)((int64_t)n * (t + 1) / nthreads)); });
	for (auto &x : th)
		x.join();
}

void copy_segments(int cap, const std::vector<Seg> &segs)
{
	size_t total = 0;
	for (auto &s : segs)
		total += s.n;
	int nthreads = (int)std::min<size_t>(cap, total / (1u << 18) + 1);
	if (nthreads <= 1) {
		for (auto &s : segs)
			memcpy(s.dst, s.src, s.n * sizeof(int32_t));
		return;
	}
	// split by element count
	const size_t per = (total + nthreads - 1) / nthreads;
	std::vector<std::thread> th;
	for (int t = 0; t < nthreads; t++) {
		th.emplace_back([&, t]() {
			size_t lo = per * t, hi = std::min(total, lo + per), pos = 0;
			for (auto &s : segs) {
				size_t b = std::max(lo, pos), e = std::min(hi, pos + s.n);
				if (b < e)
					memcpy(s.dst + (b - pos), s.src + (b - pos),
					       (e - b) * sizeof(int32_t));
				pos += s.n;
				if (pos >= hi)
					break;
			}
		});
	}
	for (auto &t : th)
		t.join();
}

// Range of everything the kernel stores or forms for a strip whose longest query / subject
// are lq / ls, in the stored coordinates of sab_wave.cuh.  The DP is
// swept over subject columns padded to the register-strip width (pad letter scores 0).
//   NW  V = H - g(r+c):  0 <= V (all-gap path) and H <= min(r,c) max(S)
//   GA  V = M - e(r+c):  2o <= V, X and Y >= 3o, M <= min(r,c) max(S)
//   SW  V = M + o, X, Y: -inf stand-ins aside, o + e <= everything <= min(r,c) max(S)
// plus the transient diagonal sum V + (S - shift) and the (o - e) steps.
void stored_range(const sab_ctx *c, int lq, int ls, int64_t *low, int64_t *high)
{
	const int64_t lsp = ((int64_t)ls + sab::kStrip - 1) / sab::kStrip * sab::kStrip;
	const int64_t gA = c->gapA, gE = c->gapE;
	const int64_t smax = std::max(c->max_s, 0), smin = std::min(c->min_s, 0);
	const int64_t top = std::min<int64_t>(lq, lsp) * smax;
	switch (c->cfg.method) {
	case SAB_NW:
		*low = smin + 2 * gA;
		*high = top - gA * (lq + lsp + 2) + smax;
		break;
	case SAB_GA:
		*low = 3 * gA + 3 * gE + smin;
		*high = top - gE * (lq + lsp + 2) + smax - gA - gE;
		break;
	default:
		*low = 2 * gA + 2 * gE + smin;
		*high = top + smax - gA;
		break;
	}
}

bool int16_safe(const sab_ctx *c, int lq, int ls)
{
	const int64_t gA = c->gapA, gE = c->gapE;
	const int64_t shift = sab::profile_shift(c->cfg.method, c->gapA, c->gapE);
	// profile entries S - shift, the per-lane gap constants and the SW threshold arithmetic
	if (c->max_s - shift > 16000 || c->min_s - shift < -16000 || -gA > 8000 || -gE > 8000)
		return false;
	int64_t low, high;
	stored_range(c, lq, ls, &low, &high);
	if (c->cfg.method == SAB_SW)
		high = std::max<int64_t>(c->max_s, 0) - gA; // wraps are caught a posteriori (re-run)
	return low >= -32000 && high <= 32000;
}

// Panels = ranges of consecutive ORIGINAL indices = the rows of one row strip.  `start` gets the
// npanels + 1 boundaries (multiples of 32: a 32-slot query group never straddles a panel).
//   panel > 0         uniform height, as asked for.
//   static shards     (sab_align_shard: every rank runs the strips a deterministic plan gives it)
//                     EQUAL-COST strips: rows [a0, a1) cost (P[a1]^2 - P[a0]^2) / 2 cells, P = prefix
//                     sum of the lengths, so the boundaries sit where P^2 crosses k / E of its final
//                     value: a tall strip at the top of the triangle, thin ones at the bottom, all
//                     with the same number of cells (and, near enough, of pairs: equal buffers).
//                     E = 2 per shard (4 for a single shard), more when a strip would exceed 2^29
//                     scores (2 GiB per strip buffer); the last strip of every shard is split into a
//                     geometric tail (below).  The launches then carry the same work however many
//                     shards there are -- uniform 6-per-shard panels left the strips of an 8-GPU job
//                     at 640 rows and the per-GPU rate 5-9 % below the 1-GPU one -- and the longest-
//                     processing-time assignment is balanced by construction.
//   dynamic           (sab_multi_*: devices pull strips from a shared cursor, costliest first)
//                     uniform height, ~6 strips per device: costs then FALL linearly along the issue
//                     order, which is what keeps the tail of a dynamically scheduled job short.
//                     With SAB_FLAG_ORDERED (ascending rows, file-backed sinks): equal-cost strips.
void panel_bounds(const std::vector<int64_t> &P /* [num + 1] prefix sums of the lengths, original order */,
		  int num, int panel, int shards_hint, bool dynamic, bool ordered, std::vector<int32_t> &start)
{
	start.clear();
	start.push_back(0);
	const int shards = std::max(1, shards_hint);
	if (panel <= 0 && dynamic && !ordered) {
		panel = (num / (6 * shards) + kPanelQuantum - 1) / kPanelQuantum * kPanelQuantum;
		panel = std::max(2 * kPanelQuantum, std::min(panel, 4096));
	}
	if (panel > 0) {
		for (int a = panel; a < num; a += panel)
			start.push_back(a);
		start.push_back(num);
		return;
	}
	// (dynamic + SAB_FLAG_ORDERED arrives here as well: ascending issue order puts the costliest
	// uniform panels LAST, which left the devices of the 200 000-sequence C5 run finishing up to a
	// strip -- 30 s -- apart; equal-cost strips, 6 per device, bound that by 1 / (6 x devices))
	const int unit = dynamic ? 6 * shards : shards > 1 ? 2 * shards : 4;
	const int64_t pairs = tri(num);
	const int E = (int)((std::max<int64_t>(1, (pairs + (1ll << 29) - 1) >> 29) + unit - 1) / unit) * unit;
	const int minh = 64; // two query groups
	// Strip costs from the bottom of the triangle upwards, in units of T = total / E.  Long
	// sequences: E equal strips (few, large launches: 5 strips per rank instead of 2 cost 4 % of the
	// kernel rate on C2, profiles/r2_ab_call2.txt).  Short sequences (copying a score costs about
	// as much as computing it: C4, 4 B per 2 025 cells, D2H + host scatter = 0.9 x the kernel time):
	// E - shards full strips, then per shard a tail of T/2, T/4, T/8, T/8 -- the write-back of a
	// rank's LAST strip overlaps with nothing, so a rank has to end on small strips; the tails sum
	// to T per shard and the longest-processing-time assignment still comes out even.
	std::vector<double> frac;
	const double cells_per_pair = (double)P[num] / num * ((double)P[num] / num);
	if (!dynamic && cells_per_pair < kTailBelowCellsPerPair) {
		frac.insert(frac.end(), (size_t)(E - shards), 1.0);
		for (double f : { 0.5, 0.25, 0.125, 0.125 })
			frac.insert(frac.end(), (size_t)shards, f);
	} else {
		frac.insert(frac.end(), (size_t)E, 1.0);
	}
	const double T2 = (double)P[num] * (double)P[num] / E; // in units of P^2 = 2 * cells
	std::vector<int32_t> cuts;
	int a1 = num;
	double hi2 = (double)P[num] * (double)P[num];
	for (size_t k = 0; k + 1 < frac.size() && a1 > 0; k++) {
		hi2 -= frac[k] * T2; // exact targets: rounding a boundary does not shift the ones above it
		int a0 = 0;
		if (hi2 > 0) {
			a0 = (int)(std::lower_bound(P.begin(), P.begin() + a1, (int64_t)std::sqrt(hi2)) - P.begin());
			a0 = (a0 + 16) / 32 * 32;
		}
		a0 = std::max(0, std::min(a0, (a1 - minh) / 32 * 32)); // a1 == num need not be a multiple of 32
		if (a0 < minh)
			a0 = 0;
		cuts.push_back(a0);
		a1 = a0;
	}
	for (int k = (int)cuts.size() - 1; k >= 0; k--)
		if (cuts[k] > 0)
			start.push_back(cuts[k]);
	start.push_back(num);
	start.erase(std::unique(start.begin(), start.end()), start.end());
}

// Longest-processing-time assignment of strips (cost ~ cells) to shards; deterministic.
void lpt_assign(const std::vector<double> &cost, int nshards, std::vector<int> &owner)
{
	const int n = (int)cost.size();
	std::vector<int> order(n);
	std::iota(order.begin(), order.end(), 0);
	std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return cost[x] > cost[y]; });
	std::vector<double> load(nshards, 0.0);
	owner.assign(n, 0);
	for (int idx : order) {
		int best = 0;
		for (int k = 1; k < nshards; k++)
			if (load[k] < load[best])
				best = k;
		load[best] += cost[idx];
		owner[idx] = best;
	}
}

// CUDA ordinals of the devices this library can drive (compute capability 10.x), in ordinal order.
std::vector<int> usable_devices()
{
	std::vector<int> out;
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		return out;
	}
	for (int d = 0; d < n; d++) {
		int major = 0;
		if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess &&
		    major == 10)
			out.push_back(d);
	}
	cudaGetLastError();
	return out;
}

bool int32_safe(const sab_ctx *c, int lq, int ls)
{
	int64_t low, high;
	stored_range(c, lq, ls, &low, &high);
	return high < (1ll << 29) && low > -(1ll << 29);
}

} // namespace

int sab_internal_device(const sab_ctx *ctx)
{
	return ctx && ctx->sm_count ? ctx->cfg.device : -1;
}

void **sab_internal_filter_slot(sab_ctx *ctx, void (*release)(void *))
{
	ctx->filter_free = release;
	return &ctx->filter_state;
}

void sab_internal_flags(const sab_ctx *ctx, int *flags)
{
	*flags = ctx ? ctx->cfg.flags : 0;
}

void sab_internal_filter_stats(sab_ctx *ctx, double ms, int64_t matches)
{
	ctx->stats.filter_ms = ms;
	ctx->stats.filter_matches = matches;
}

int sab_internal_fail(sab_ctx *ctx, int code, const char *fmt, ...)
{
	char b[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(b, sizeof b, fmt, ap);
	va_end(ap);
	if (ctx)
		ctx->err = b;
	return code;
}

extern "C" {

int sab_abi_version(void)
{
	return SAB_ABI_VERSION;
}

int sab_device_count(void)
{
	return (int)usable_devices().size();
}

int sab_device_ordinal(int index)
{
	const std::vector<int> dev = usable_devices();
	return index >= 0 && index < (int)dev.size() ? dev[index] : SAB_ERR_NODEVICE;
}

int sab_device_memory(int device, size_t *free_bytes, size_t *total_bytes)
{
	int cur = 0;
	if (cudaGetDevice(&cur) != cudaSuccess || cudaSetDevice(device) != cudaSuccess) {
		cudaGetLastError();
		return SAB_ERR_NODEVICE;
	}
	size_t f = 0, t = 0;
	cudaError_t e = cudaMemGetInfo(&f, &t);
	cudaSetDevice(cur);
	if (e != cudaSuccess)
		return SAB_ERR_CUDA;
	if (free_bytes)
		*free_bytes = f;
	if (total_bytes)
		*total_bytes = t;
	return SAB_OK;
}

const char *sab_last_error(const sab_ctx *ctx)
{
	return ctx ? ctx->err.c_str() : "null context";
}

int sab_create(sab_ctx **out, const sab_config *cfg)
{
	if (!out || !cfg)
		return SAB_ERR_ARG;
	*out = nullptr;
	sab_ctx *ctx = new (std::nothrow) sab_ctx;
	if (!ctx)
		return SAB_ERR_NOMEM;
	*out = ctx; // returned even on failure so sab_last_error works; caller destroys
	ctx->cfg = *cfg;
	if (cfg->method < SAB_NW || cfg->method > SAB_SW)
		return fail(ctx, SAB_ERR_ARG, "unknown method %d", cfg->method);
	if (cfg->gap_pen > 0 || cfg->gap_open > 0 || cfg->gap_ext > 0)
		return fail(ctx, SAB_ERR_ARG, "gap values must be stored <= 0 (align.c:127)");
	if (cfg->panel < 0 || cfg->panel % kPanelQuantum)
		return fail(ctx, SAB_ERR_ARG, "panel must be a multiple of %d", kPanelQuantum);
	// Gotoh with |extend| >= |open| IS Needleman-Wunsch with g = open: M >= X in every cell
	// (ga.c:61-63), so X = max(Mleft + o, Xleft + e) always takes its first term when e <= o,
	// and the border recurrence (ga.c:26-37) gives M[0][c] = c*o, not o + (c-1)e.  The strip
	// kernel's constant GA border only holds for |e| <= |o|, so such configurations (legal on
	// the CLI: align.c puts no constraint on -s versus -e; equal values are what validate_ga,
	// ga.c:70-88, offers to switch) run as NW -- exact, and a third of the alu work.
	if (cfg->method == SAB_GA && cfg->gap_ext <= cfg->gap_open) {
		ctx->cfg.method = SAB_NW;
		ctx->cfg.gap_pen = cfg->gap_open;
		cfg = &ctx->cfg;
	}
	ctx->gapA = cfg->method == SAB_NW ? cfg->gap_pen : cfg->gap_open;
	ctx->gapE = cfg->method == SAB_NW ? 0 : cfg->gap_ext;
	// symmetric tables only (all 67 reference tables are; the kernels swap the
	// operand order freely between nw.c:27 and ga.c:46)
	int mn = INT32_MAX, mx = INT32_MIN;
	for (int a = 0; a < 24; a++)
		for (int b = 0; b < 24; b++) {
			const int v = cfg->sub_mat[a * 24 + b];
			if (v != cfg->sub_mat[b * 24 + a])
				return fail(ctx, SAB_ERR_ARG, "substitution matrix is not symmetric");
			mn = std::min(mn, v);
			mx = std::max(mx, v);
		}
	ctx->min_s = mn;
	ctx->max_s = mx;
	if ((int64_t)mx - mn > (1 << 20) || -(int64_t)ctx->gapA > (1 << 28) ||
	    -(int64_t)ctx->gapE > (1 << 28))
		return fail(ctx, SAB_ERR_ARG, "scores out of the supported int32 range");
	for (int a = 0; a < sab::kLetters; a++)
		for (int b = 0; b < sab::kLetters; b++) {
			const int v = (a < 24 && b < 24) ? cfg->sub_mat[a * 24 + b] : 0;
			ctx->sub25[a * sab::kLetters + b] =
				v - sab::profile_shift(cfg->method, ctx->gapA, ctx->gapE);
		}

	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
		cudaGetLastError();
		return fail(ctx, SAB_ERR_NODEVICE, "no CUDA device (this library has no CPU fallback)");
	}
	if (cfg->device < 0 || cfg->device >= ndev)
		return fail(ctx, SAB_ERR_NODEVICE, "device %d out of range (%d present)", cfg->device,
			    ndev);
	int major = 0, minor = 0, sms = 0;
	CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cfg->device));
	CU(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, cfg->device));
	CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device));
	if (major != 10)
		return fail(ctx, SAB_ERR_NODEVICE,
			    "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device,
			    major, minor);
	CU(cudaSetDevice(cfg->device));
	ctx->sm_count = sms;
	CU(cudaMalloc(&ctx->d_sub25, sizeof ctx->sub25));
	CU(cudaMemcpy(ctx->d_sub25, ctx->sub25, sizeof ctx->sub25, cudaMemcpyHostToDevice));
	// Kernel variants are set up on first use (wave_variant): a job runs one or two of the twelve.
	// The pinned bounce buffers and the SW re-run list are allocated on first use as well.
	for (int i = 0; i < 2; i++) {
		StripBuf &b = ctx->buf[i];
		CU(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
		CU(cudaEventCreate(&b.ev_start));
		CU(cudaEventCreate(&b.ev_done));
		CU(cudaMalloc(&b.counter, kCounterWords * sizeof(unsigned)));
		CU(cudaEventCreateWithFlags(&b.ev_fork, cudaEventDisableTiming));
		CU(cudaEventCreateWithFlags(&ctx->ev_copy[i], cudaEventDisableTiming));
	}
	CU(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
	CU(cudaEventCreate(&ctx->ev_span[0]));
	CU(cudaEventCreate(&ctx->ev_span[1]));
	return SAB_OK;
}

void sab_destroy(sab_ctx *ctx)
{
	if (!ctx)
		return;
	if (ctx->sm_count) {
		cudaSetDevice(ctx->cfg.device);
		cudaDeviceSynchronize();
		if (ctx->filter_state && ctx->filter_free)
			ctx->filter_free(ctx->filter_state);
		sab_module_release(ctx->modules);
		free_seqs(ctx);
		free_bufs(ctx);
		cudaFree(ctx->d_sub25);
		for (int i = 0; i < 2; i++) {
			StripBuf &b = ctx->buf[i];
			cudaFree(b.counter);
			cudaFree(b.rerun_list);
			if (b.stream)
				cudaStreamDestroy(b.stream);
			if (b.ev_start)
				cudaEventDestroy(b.ev_start);
			if (b.ev_done)
				cudaEventDestroy(b.ev_done);
			if (b.ev_fork)
				cudaEventDestroy(b.ev_fork);
			for (auto s : b.side)
				if (s)
					cudaStreamDestroy(s);
			for (auto e : b.ev_join)
				if (e)
					cudaEventDestroy(e);
			if (ctx->stage[i])
				cudaFreeHost(ctx->stage[i]);
			if (ctx->ev_copy[i])
				cudaEventDestroy(ctx->ev_copy[i]);
		}
		if (ctx->copy_stream)
			cudaStreamDestroy(ctx->copy_stream);
		for (auto e : ctx->ev_span)
			if (e)
				cudaEventDestroy(e);
		cudaGetLastError();
	}
	delete ctx;
}

} // extern "C"

namespace {

// The sequence set in its two device layouts, packed once on the host;
// sab_multi_upload hands the same object to every device.
struct PackedSet {
	int32_t num = 0, max_len = 0;
	std::vector<int32_t> slot_len, slot_orig, panel_maxlen, panel_start;
	std::vector<int64_t> col_off, grp_off, len_prefix, pad16_prefix, pad32_prefix;
	std::vector<uint32_t> col_words; // subject letters, col_bits bits each; every slot starts on a word
	int col_bits = 5;
	uint32_t col_dec = 0; // 2-bit code -> letter index, one byte per code
	std::vector<uint32_t> rowpack;
};

// Argument checks + length binning + packing.  `ctx` supplies the configuration (LUT, panel,
// shards hint, ranges) and receives the error text.
int pack_set(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num, int32_t max, PackedSet &ps)
{
	if (!seqs || !meta || num < 2)
		return fail(ctx, SAB_ERR_ARG, "need at least 2 sequences (SEQ_N_MIN, align.h:21)");
	int32_t real_max = 0;
	int64_t letters = 0;
	for (int32_t k = 0; k < num; k++) {
		if (meta[k].len < 1 || meta[k].off < 0)
			return fail(ctx, SAB_ERR_ARG, "sequence %d has bad length/offset", k);
		real_max = std::max(real_max, meta[k].len);
		letters += meta[k].len;
	}
	if (max > 0 && max < real_max)
		return fail(ctx, SAB_ERR_ARG, "max (%d) is below the longest sequence (%d)", max,
			    real_max);
	if (!int32_safe(ctx, real_max, real_max))
		return fail(ctx, SAB_ERR_ARG, "lengths x scores exceed the int32 score range");

	// Panels: contiguous ranges of ORIGINAL indices, so a row strip of the output is
	// one contiguous run of the packed-triangular matrix.
	{
		std::vector<int64_t> P((size_t)num + 1, 0);
		for (int32_t k = 0; k < num; k++)
			P[k + 1] = P[k] + meta[k].len;
		panel_bounds(P, num, ctx->cfg.panel, ctx->cfg.shards_hint, ctx->cfg.flags & SAB_FLAG_INTERNAL_DYNAMIC,
			     ctx->cfg.flags & SAB_FLAG_ORDERED, ps.panel_start);
	}
	const int npanels = (int)ps.panel_start.size() - 1;

	// Length binning: stable sort by length inside each panel -> slots.
	std::vector<int32_t> &slot_orig = ps.slot_orig, &slot_len = ps.slot_len;
	slot_orig.resize(num);
	slot_len.resize(num);
	std::iota(slot_orig.begin(), slot_orig.end(), 0);
	ps.panel_maxlen.assign(npanels, 0);
	for (int p = 0; p < npanels; p++) {
		auto b = slot_orig.begin() + ps.panel_start[p];
		auto e = slot_orig.begin() + ps.panel_start[p + 1];
		std::stable_sort(b, e, [&](int32_t x, int32_t y) { return meta[x].len < meta[y].len; });
		ps.panel_maxlen[p] = meta[*(e - 1)].len;
	}
	ps.len_prefix.resize(num + 1);
	ps.pad16_prefix.resize(num + 1);
	ps.pad32_prefix.resize(num + 1);
	int64_t off = 0, off16 = 0, off32 = 0;
	for (int32_t s = 0; s < num; s++) {
		slot_len[s] = meta[slot_orig[s]].len;
		ps.len_prefix[s] = off;
		ps.pad16_prefix[s] = off16;
		ps.pad32_prefix[s] = off32;
		off += slot_len[s];
		off16 += (slot_len[s] + 15) / 16 * 16;
		off32 += (slot_len[s] + 31) / 32 * 32;
	}
	ps.len_prefix[num] = off;
	ps.pad16_prefix[num] = off16;
	ps.pad32_prefix[num] = off32;

	// Alphabet check; which of the 24 letter indices occur at all decides the packing below.
	std::atomic<int32_t> bad_seq{ -1 };
	std::atomic<uint32_t> used{ 0 };
	parallel_ranges(host_thread_cap(ctx), num, letters, [&](int32_t s0, int32_t s1) {
		uint32_t mine = 0;
		for (int32_t s = s0; s < s1; s++) {
			const uint8_t *src = seqs + meta[slot_orig[s]].off;
			for (int32_t k = 0; k < slot_len[s]; k++) {
				const uint8_t ch = src[k];
				const int32_t idx = ch < 128 ? ctx->cfg.seq_lut[ch] : -1;
				if (idx < 0 || idx >= 24) {
					bad_seq.store(slot_orig[s]);
					return;
				}
				mine |= 1u << idx;
			}
		}
		used.fetch_or(mine);
	});
	if (bad_seq.load() >= 0)
		return fail(ctx, SAB_ERR_ARG, "sequence %d has a letter outside the matrix alphabet",
			    bad_seq.load());
	// Subject (DP column) layout: a bit stream per slot, 2 bits per letter when the set uses at most
	// four letters of the matrix alphabet (DNA reads: ACGT), else 5 bits (protein, IUPAC); every slot
	// starts on a 32-bit word, so the packing threads never share a word.  Read once per strip when
	// its profile is staged (sab_wave.cuh: build) and by the int32 re-run kernel.
	uint8_t code_of[24] = { 0 };
	{
		const uint32_t u = used.load();
		int distinct = 0;
		for (int a = 0; a < 24; a++)
			distinct += (u >> a) & 1;
		ps.col_bits = distinct <= 4 ? 2 : 5;
		ps.col_dec = 0;
		int next = 0;
		for (int a = 0; a < 24; a++) {
			code_of[a] = (uint8_t)a;
			if (ps.col_bits == 2 && ((u >> a) & 1)) {
				code_of[a] = (uint8_t)next;
				ps.col_dec |= (uint32_t)a << (8 * next);
				next++;
			}
		}
	}
	const int bits = ps.col_bits;
	std::vector<int64_t> &col_off = ps.col_off; // WORD offset of the slot in col_words
	col_off.resize(num);
	int64_t words = 0;
	for (int32_t s = 0; s < num; s++) {
		col_off[s] = words;
		words += ((int64_t)slot_len[s] * bits + 31) / 32;
	}
	std::vector<uint32_t> &col_words = ps.col_words;
	col_words.assign((size_t)words + 1, 0u);
	parallel_ranges(host_thread_cap(ctx), num, letters, [&](int32_t s0, int32_t s1) {
		for (int32_t s = s0; s < s1; s++) {
			const uint8_t *src = seqs + meta[slot_orig[s]].off;
			uint32_t *dst = col_words.data() + col_off[s];
			for (int32_t k = 0; k < slot_len[s]; k++) {
				const uint32_t code = code_of[ctx->cfg.seq_lut[src[k]]];
				const int64_t b = (int64_t)k * bits;
				const int sh = (int)(b & 31);
				dst[b >> 5] |= code << sh;
				if (sh + bits > 32)
					dst[(b >> 5) + 1] |= code >> (32 - sh);
			}
		}
	});
	// Query layout: 32-slot groups, 4 residues per word (stored as byte offsets
	// letter*4 into a profile column), lane-interleaved for coalesced loads.
	const int ngroups = (num + 31) / 32;
	std::vector<int64_t> &grp_off = ps.grp_off;
	grp_off.resize(ngroups);
	int64_t rows4 = 0;
	for (int g = 0; g < ngroups; g++) {
		grp_off[g] = rows4;
		const int last = std::min(num, (g + 1) * 32) - 1;
		int gmax = 0;
		for (int s = g * 32; s <= last; s++)
			gmax = std::max(gmax, slot_len[s]);
		rows4 += (gmax + 3) / 4 + 2; // + prefetch slack
	}
	// rows past a sequence's end hold the pad letter (24): SW sweeps them (sab_wave.cuh)
	std::vector<uint32_t> &rowpack = ps.rowpack;
	rowpack.assign((size_t)rows4 * 32, 0x60606060u);
	parallel_ranges(host_thread_cap(ctx), num, letters, [&](int32_t s0, int32_t s1) { // every slot owns its own words
		for (int32_t s = s0; s < s1; s++) {
			const uint8_t *src = seqs + meta[slot_orig[s]].off;
			uint32_t *dst = rowpack.data() + grp_off[s >> 5] * 32 + (s & 31);
			for (int32_t r = 0; r < slot_len[s]; r++) {
				uint32_t &wd = dst[(size_t)(r >> 2) * 32];
				const int sh = 8 * (r & 3);
				wd = (wd & ~(0xffu << sh)) | ((uint32_t)(ctx->cfg.seq_lut[src[r]] * 4) << sh);
			}
		}
	});
	ps.num = num;
	ps.max_len = real_max;
	return SAB_OK;
}

// Copies a packed set to the context's device; `keep` = the host tables may be moved out of `ps`.
int install_set(sab_ctx *ctx, PackedSet &ps, bool keep)
{
	if (!ctx->sm_count)
		return fail(ctx, SAB_ERR_STATE, "context has no device");
	CU(cudaSetDevice(ctx->cfg.device));
	ctx->uploaded = false; // device buffers are kept and reused when large enough
	const int32_t num = ps.num;
	const int ngroups = (int)ps.grp_off.size();
	// cudaFree / cudaMalloc per upload cost 0.1-0.25 s every few calls (measured): grow-only reuse
	{
		void **ptr[6] = { (void **)&ctx->d_slot_len, (void **)&ctx->d_slot_orig, (void **)&ctx->d_col_off,
				  (void **)&ctx->d_grp_off, (void **)&ctx->d_col_words, (void **)&ctx->d_rowpack };
		const size_t need[6] = { sizeof(int32_t) * num, sizeof(int32_t) * num, sizeof(int64_t) * num,
					 sizeof(int64_t) * ngroups, ps.col_words.size() * sizeof(uint32_t),
					 ps.rowpack.size() * sizeof(uint32_t) };
		for (int i = 0; i < 6; i++) {
			const int rc = ensure(ctx, ptr[i], &ctx->cap_seqs[i], need[i], 1);
			if (rc)
				return rc;
		}
	}
	CU(cudaMemcpy(ctx->d_slot_len, ps.slot_len.data(), sizeof(int32_t) * num, cudaMemcpyHostToDevice));
	CU(cudaMemcpy(ctx->d_slot_orig, ps.slot_orig.data(), sizeof(int32_t) * num, cudaMemcpyHostToDevice));
	CU(cudaMemcpy(ctx->d_col_off, ps.col_off.data(), sizeof(int64_t) * num, cudaMemcpyHostToDevice));
	CU(cudaMemcpy(ctx->d_grp_off, ps.grp_off.data(), sizeof(int64_t) * ngroups, cudaMemcpyHostToDevice));
	CU(cudaMemcpy(ctx->d_col_words, ps.col_words.data(), ps.col_words.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
	ctx->col_bits = ps.col_bits;
	ctx->col_dec = ps.col_dec;
	CU(cudaMemcpy(ctx->d_rowpack, ps.rowpack.data(), ps.rowpack.size() * sizeof(uint32_t),
		      cudaMemcpyHostToDevice));
	ctx->stats.h2d_bytes = (int64_t)(sizeof(int32_t) * 2 * num + sizeof(int64_t) * (num + ngroups) +
					 ps.col_words.size() * sizeof(uint32_t) + ps.rowpack.size() * sizeof(uint32_t));
	ctx->num = num;
	ctx->max_len = ps.max_len;
	if (keep) {
		ctx->slot_len = ps.slot_len;
		ctx->slot_orig = ps.slot_orig;
		ctx->len_prefix = ps.len_prefix;
		ctx->pad16_prefix = ps.pad16_prefix;
		ctx->pad32_prefix = ps.pad32_prefix;
		ctx->panel_maxlen = ps.panel_maxlen;
		ctx->panel_start = ps.panel_start;
	} else {
		ctx->slot_len.swap(ps.slot_len);
		ctx->slot_orig.swap(ps.slot_orig);
		ctx->len_prefix.swap(ps.len_prefix);
		ctx->pad16_prefix.swap(ps.pad16_prefix);
		ctx->pad32_prefix.swap(ps.pad32_prefix);
		ctx->panel_maxlen.swap(ps.panel_maxlen);
		ctx->panel_start.swap(ps.panel_start);
	}
	ctx->sw_prefer_int32 = false;
	ctx->uploaded = true;
	return SAB_OK;
}

} // namespace

extern "C" {

int sab_upload(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num, int32_t max)
{
	if (!ctx)
		return SAB_ERR_ARG;
	if (!ctx->sm_count)
		return fail(ctx, SAB_ERR_STATE, "context has no device");
	const double t0 = now_ms();
	PackedSet ps;
	int rc = pack_set(ctx, seqs, meta, num, max, ps);
	if (rc == SAB_OK)
		rc = install_set(ctx, ps, false);
	if (rc)
		return rc;
	ctx->stats.upload_ms = now_ms() - t0;
	return SAB_OK;
}

// Host-only self-test of the upload packing (no device needed): packs the set as sab_upload would
// and reads every letter back from the two device layouts with the kernels' own addressing --
// the bit stream of the subject layout (sab_kernels.cuh: subject_letter) and the lane-interleaved
// byte-offset words of the query layout.  Returns the number of letters that differ (0 = good),
// < 0 on argument errors; *bits_per_letter reports the subject packing chosen (2 or 5).
int sab_selftest_packing(const sab_config *cfg, const uint8_t *seqs, const sab_meta *meta, int32_t num,
			 int32_t *bits_per_letter)
{
	if (!cfg || !seqs || !meta)
		return SAB_ERR_ARG;
	sab_ctx tmp;
	tmp.cfg = *cfg;
	tmp.gapA = cfg->method == SAB_NW ? cfg->gap_pen : cfg->gap_open;
	tmp.gapE = cfg->method == SAB_NW ? 0 : cfg->gap_ext;
	tmp.max_s = 1;
	PackedSet ps;
	const int rc = pack_set(&tmp, seqs, meta, num, 0, ps);
	if (rc)
		return rc;
	if (bits_per_letter)
		*bits_per_letter = ps.col_bits;
	int64_t bad = 0;
	for (int32_t s = 0; s < num; s++) {
		const uint8_t *src = seqs + meta[ps.slot_orig[s]].off;
		const uint32_t *w = ps.col_words.data() + ps.col_off[s];
		const uint32_t *rp = ps.rowpack.data() + ps.grp_off[s >> 5] * 32 + (s & 31);
		if (ps.slot_len[s] != meta[ps.slot_orig[s]].len)
			bad++;
		for (int32_t k = 0; k < ps.slot_len[s]; k++) {
			const int want = cfg->seq_lut[src[k]];
			const int b = k * ps.col_bits, sh = b & 31;
			uint32_t v = w[b >> 5] >> sh;
			if (sh + ps.col_bits > 32)
				v |= w[(b >> 5) + 1] << (32 - sh);
			const int got = ps.col_bits == 2 ? (int)((ps.col_dec >> (8 * (v & 3u))) & 0xffu) : (int)(v & 31u);
			const int rowgot = (int)((rp[(size_t)(k >> 2) * 32] >> (8 * (k & 3))) & 0xffu) / 4;
			bad += (got != want) + (rowgot != want);
		}
	}
	return (int)std::min<int64_t>(bad, INT32_MAX);
}

// Strip bookkeeping ------------------------------------------------------------------------

namespace {

struct Strip {
	int panel; // index
	int a0, an; // original index range == slot range
	double cost; // sum len_q * sum len_s (approx. cells)
	bool packed;
	int widx_fixed; // index into kWidths when the whole strip runs at one width (int32 strips, the
			// STRIP32/STRIP16 flags), -1 = chosen at launch time (choose_widths)
	int lq, ls; // longest query / subject
	double mean_ls; // mean subject length (strip width and chunk size heuristics)
};


// Row strips (one per panel), their cost, and the longest-processing-time assignment
// of strips to shards (deterministic: every rank computes the same plan).
void plan_strips(const sab_ctx *ctx, int nshards, std::vector<Strip> &strips, std::vector<int> &owner)
{
	const int npanels = (int)ctx->panel_start.size() - 1;
	const bool force32 = ctx->cfg.flags & SAB_FLAG_FORCE_INT32;
	const bool force16 = ctx->cfg.flags & SAB_FLAG_FORCE_INT16;
	strips.assign(npanels, Strip{});
	int run_max = 0;
	for (int p = 0; p < npanels; p++) {
		Strip &s = strips[p];
		s.panel = p;
		s.a0 = ctx->panel_start[p];
		s.an = ctx->panel_start[p + 1] - s.a0;
		run_max = std::max(run_max, ctx->panel_maxlen[p]);
		s.lq = ctx->panel_maxlen[p];
		s.ls = run_max;
		const double lq_sum = (double)(ctx->len_prefix[s.a0 + s.an] - ctx->len_prefix[s.a0]);
		const double ls_sum = (double)ctx->len_prefix[s.a0] + 0.5 * lq_sum;
		s.cost = lq_sum * ls_sum;
		s.packed = !force32 && (force16 || int16_safe(ctx, s.lq, s.ls));
		// Strip width: packed strips choose it at launch time (choose_widths).
		// int32 strips exist at the widest and the narrowest width only: the subjects of the strip
		// (every slot below its last query) padded to 16 or to 32 DP columns, weighted by what a
		// column costs at either width.
		const int a1 = s.a0 + s.an;
		const double mean_ls = (double)ctx->len_prefix[a1] / a1;
		s.mean_ls = mean_ls;
		const double cols16 = (double)ctx->pad16_prefix[a1], cols32 = (double)ctx->pad32_prefix[a1];
		if (ctx->cfg.flags & SAB_FLAG_STRIP16)
			s.widx_fixed = kNarrowest;
		else if (ctx->cfg.flags & SAB_FLAG_STRIP32)
			s.widx_fixed = kWidest;
		else if (!s.packed)
			s.widx_fixed = cols16 * column_cost(16) < cols32 * column_cost(32) ? kNarrowest : kWidest;
		else
			s.widx_fixed = -1;
	}
	std::vector<double> cost(npanels);
	for (int p = 0; p < npanels; p++)
		cost[p] = strips[p].cost;
	lpt_assign(cost, nshards, owner);
}

// Copies the finished strip (device) into the caller's matrix through the two pinned
// bounce buffers.  tri: the strip IS matrix[tri(a0) .. tri(a0+an)).  full: lower-row
// parts from blk, mirrored parts from blkT (output_fill, src/io/output.c:72-78).
int drain_strip(sab_ctx *ctx, const Strip &st, StripBuf &b, int32_t *matrix, int triangular)
{
	const int64_t N = ctx->num;
	const int64_t a0 = st.a0, a1 = st.a0 + st.an;
	const size_t stage_elems = kStageBytes / sizeof(int32_t);
	for (int i = 0; i < 2; i++) // first write-back of this context (-W never gets here)
		if (!ctx->stage[i])
			CU(cudaHostAlloc(&ctx->stage[i], kStageBytes, cudaHostAllocDefault));
	// pass 0 = lower rows (blk), pass 1 = mirror (blkT); two chunks in flight.
	int slot = 0;
	bool pending[2] = { false, false };
	std::vector<Seg> pend_segs[2];
	auto flush_slot = [&](int s) -> int {
		if (!pending[s])
			return SAB_OK;
		CU(cudaEventSynchronize(ctx->ev_copy[s]));
		copy_segments(host_thread_cap(ctx), pend_segs[s]);
		pending[s] = false;
		return SAB_OK;
	};
	auto submit = [&](const int32_t *dev, size_t n, std::vector<Seg> &&segs) -> int {
		int rc = flush_slot(slot);
		if (rc)
			return rc;
		CU(cudaMemcpyAsync(ctx->stage[slot], dev, n * sizeof(int32_t), cudaMemcpyDeviceToHost,
				   ctx->copy_stream));
		CU(cudaEventRecord(ctx->ev_copy[slot], ctx->copy_stream));
		ctx->stats.d2h_bytes += (int64_t)(n * sizeof(int32_t));
		pend_segs[slot] = std::move(segs);
		pending[slot] = true;
		slot ^= 1;
		return SAB_OK;
	};

	// lower rows
	{
		const int64_t total = tri(a1) - tri(a0);
		int64_t e = 0;
		int64_t row = a0; // row containing element e
		while (e < total) {
			const size_t n = (size_t)std::min<int64_t>(stage_elems, total - e);
			const int32_t *stage = (const int32_t *)ctx->stage[slot];
			std::vector<Seg> segs;
			if (triangular) {
				segs.push_back({ matrix + tri(a0) + e, stage, n });
			} else {
				// rows overlapping [e, e+n)
				int64_t pos = e;
				while (pos < e + (int64_t)n) {
					while (tri(row + 1) - tri(a0) <= pos)
						row++;
					const int64_t rbeg = tri(row) - tri(a0);
					const int64_t col = pos - rbeg;
					const int64_t cnt = std::min<int64_t>(row - col, e + n - pos);
					if (cnt > 0)
						segs.push_back({ matrix + row * N + col, stage + (pos - e),
								 (size_t)cnt });
					pos += cnt > 0 ? cnt : 1;
				}
			}
			int rc = submit(b.blk + e, n, std::move(segs));
			if (rc)
				return rc;
			e += n;
		}
	}
	if (!triangular) {
		// mirror: blkT[lo * an + (hi - a0)], lo in [0, a1); valid columns hi > lo
		const int64_t an = st.an;
		const int64_t rows_per = std::max<int64_t>(1, stage_elems / an);
		for (int64_t lo0 = 0; lo0 < a1 - 1; lo0 += rows_per) {
			const int64_t lo1 = std::min<int64_t>(a1 - 1, lo0 + rows_per);
			const int32_t *stage = (const int32_t *)ctx->stage[slot];
			std::vector<Seg> segs;
			for (int64_t lo = lo0; lo < lo1; lo++) {
				const int64_t h0 = std::max<int64_t>(a0, lo + 1);
				segs.push_back({ matrix + lo * N + h0, stage + (lo - lo0) * an + (h0 - a0),
						 (size_t)(a1 - h0) });
			}
			int rc = submit(b.blkT + lo0 * an, (size_t)((lo1 - lo0) * an), std::move(segs));
			if (rc)
				return rc;
		}
	}
	int rc = flush_slot(0);
	if (rc)
		return rc;
	return flush_slot(1);
}

} // namespace

namespace {

// Where a device gets its next row strip from: `order` lists strip indices in issue order and
// `cursor` hands them out -- private to one shard of the static plan (sab_align_shard), or shared
// by the devices of one process (sab_multi_align: the "shared atomic tile cursor" of SURVEY 8e, so
// a device that finishes early simply takes more strips).
struct StripSource {
	const std::vector<int> *order;
	std::atomic<int> *cursor;
	int next()
	{
		const int k = cursor->fetch_add(1);
		return k < (int)order->size() ? (*order)[k] : -1;
	}
};

// Issue order of a set of strips: costliest first (the tail of the job is then made of the cheap
// ones), or ascending rows under SAB_FLAG_ORDERED (ascending offsets of a file-backed matrix).
void issue_order(const sab_ctx *ctx, const std::vector<Strip> &strips, std::vector<int> &order)
{
	if (ctx->cfg.flags & SAB_FLAG_ORDERED) {
		std::sort(order.begin(), order.end());
		return;
	}
	std::stable_sort(order.begin(), order.end(),
			 [&](int x, int y) { return strips[x].cost > strips[y].cost; });
}

// Register-strip width of a packed strip: the one that minimises sum(ceil(pair length / K) * K) *
// column_cost(K) over its subject pairs (C3: len 150 = 5 x 30 -> K = 30, +7 % over K = 32).
// Under SAB_FLAG_PER_CHUNK_WIDTHS every subject chunk gets its own width instead (slots are
// length-sorted inside a panel, so the 16..128 subjects of a chunk have near-equal lengths) and the
// strip is launched once per width in use: measured SLOWER than one width on C2/C4/C5
// (profiles/r2_ab_call2.txt: 8.13 against 8.87 TCUPS on C2) -- a launch per width leaves each with
// too few work items for 592 persistent blocks, and kernels of different widths sharing an SM
// evict each other's 64 KB of straight-line code from the instruction cache.
// Out: ids = chunk indices grouped by width (empty: one width), first[w]..first[w+1] the group of kWidths[w].
void choose_widths(const sab_ctx *ctx, int a1, int chunk, int nsc, bool per_chunk, std::vector<int> &ids, int *first)
{
	std::vector<uint8_t> best((size_t)nsc);
	std::vector<float> cost((size_t)nsc * kNumWidths);
	int count[kNumWidths] = { 0 };
	bool enabled[kNumWidths];
	for (int w = 0; w < kNumWidths; w++)
		enabled[w] = true;
	const int32_t *len = ctx->slot_len.data();
	for (int k = 0; k < nsc; k++) {
		const int s0 = k * chunk, s1 = std::min(a1, s0 + chunk);
		double c[kNumWidths] = { 0 };
		for (int s = s0; s < s1; s += 2) {
			const int l = s + 1 < s1 ? std::max(len[s], len[s + 1]) : len[s];
			for (int w = 0; w < kNumWidths; w++) {
				const int kk = kWidths[w].k;
				c[w] += (double)((l + kk - 1) / kk * kk);
			}
		}
		for (int w = 0; w < kNumWidths; w++)
			cost[(size_t)k * kNumWidths + w] = (float)(c[w] * column_cost(kWidths[w].k));
	}
	if (!per_chunk) { // one width for the whole strip: the cheapest in total
		int bw = 0;
		double bsum = 0;
		for (int w = 0; w < kNumWidths; w++) {
			double sum = 0;
			for (int k = 0; k < nsc; k++)
				sum += cost[(size_t)k * kNumWidths + w];
			if (w == 0 || sum < bsum) {
				bw = w;
				bsum = sum;
			}
		}
		for (int w = 0; w <= kNumWidths; w++)
			first[w] = w > bw ? nsc : 0;
		ids.clear();
		return;
	}
	for (int pass = 0; pass < 2; pass++) {
		for (int w = 0; w < kNumWidths; w++)
			count[w] = 0;
		for (int k = 0; k < nsc; k++) {
			int bw = -1;
			for (int w = 0; w < kNumWidths; w++)
				if (enabled[w] && (bw < 0 || cost[(size_t)k * kNumWidths + w] < cost[(size_t)k * kNumWidths + bw]))
					bw = w;
			best[k] = (uint8_t)bw;
			count[bw]++;
		}
		if (pass == 0) {
			const int floor_ = std::max(4, nsc / 50);
			int top = 0;
			for (int w = 1; w < kNumWidths; w++)
				if (count[w] > count[top])
					top = w;
			for (int w = 0; w < kNumWidths; w++)
				enabled[w] = w == top || count[w] >= floor_;
		}
	}
	first[0] = 0;
	for (int w = 0; w < kNumWidths; w++)
		first[w + 1] = first[w] + count[w];
	ids.resize((size_t)nsc);
	int fill[kNumWidths];
	for (int w = 0; w < kNumWidths; w++)
		fill[w] = first[w];
	for (int k = 0; k < nsc; k++)
		ids[fill[best[k]]++] = k;
}

int align_strips(sab_ctx *ctx, const std::vector<Strip> &strips, StripSource &src, int64_t pairs_total,
		 int32_t *matrix, int triangular, sab_progress_fn progress, void *user)
{
	CU(cudaSetDevice(ctx->cfg.device));
	const double t0 = now_ms();
	const int num = ctx->num;
	const int method = ctx->cfg.method;
	std::vector<int> mine; // the strips this call ended up computing

	sab_stats &S = ctx->stats;
	const double up_ms = S.upload_ms, flt_ms = S.filter_ms;
	const int64_t h2d = S.h2d_bytes, flt_m = S.filter_matches;
	memset(&S, 0, sizeof S);
	S.upload_ms = up_ms;
	S.h2d_bytes = h2d;
	S.filter_ms = flt_ms;
	S.filter_matches = flt_m;

	DeviceSeqs dseqs{ ctx->d_slot_len, ctx->d_slot_orig, ctx->d_col_off, ctx->d_col_words,
			  ctx->d_grp_off, ctx->d_rowpack,   ctx->col_bits,   ctx->col_dec };
	const int bnd_rows = (ctx->max_len + 8 + 3) / 4 * 4;
	int64_t pairs_done = 0;
	std::vector<std::pair<float, float>> spans; // kernel intervals on the device timeline
	const bool test_cap = ctx->cfg.flags & SAB_FLAG_TEST_RERUN_CAP;
	const unsigned rerun_cap = test_cap ? 4u : kRerunCap;

	auto launch = [&](const Strip &st_in, StripBuf &b) -> int {
		Strip st = st_in;
		const bool packed = st.packed;
		const int a1 = st.a0 + st.an;
		// a packed strip that is swept (again) in int32 -- re-run list overflow, sw_prefer_int32 --
		// arrives without a fixed width: the int32 kernel exists at the widest and narrowest only
		if (!packed && (st.widx_fixed < 0 || (st.widx_fixed != kWidest && st.widx_fixed != kNarrowest)))
			st.widx_fixed = (double)ctx->pad16_prefix[a1] * column_cost(16) <
							(double)ctx->pad32_prefix[a1] * column_cost(32) ?
						kNarrowest :
						kWidest;
		const size_t blk_elems = (size_t)(tri(a1) - tri(st.a0));
		int rc = ensure(ctx, (void **)&b.blk, &b.blk_cap, std::max<size_t>(blk_elems, 1), sizeof(int32_t));
		if (rc)
			return rc;
		if (!triangular && matrix) {
			rc = ensure(ctx, (void **)&b.blkT, &b.blkT_cap, (size_t)a1 * st.an, sizeof(int32_t));
			if (rc)
				return rc;
		}
		if (method == SAB_SW && packed && !b.rerun_list)
			CU(cudaMalloc(&b.rerun_list, sizeof(unsigned long long) * kRerunCap));
		// Work item = 32 queries x `chunk` subjects.  An item pays 3 steps of pipeline fill, so
		// short subjects get more of them per item: aim at >= 24 rounds (4 strips each).
		const int est_strips = std::max(1, (int)((st.mean_ls + sab::kStrip - 1) / sab::kStrip));
		const int mult = std::min(8, std::max(1, (24 + 2 * est_strips - 1) / (2 * est_strips)));
		const int chunk = (packed ? 16 : 8) * mult;
		const int nsc = (a1 + chunk - 1) / chunk;
		// which chunks run at which width: lists[w] = [first[w], first[w + 1]) of b.chunk_host
		int first[kNumWidths + 1] = { 0 };
		if (st.widx_fixed >= 0) {
			for (int w = 0; w <= kNumWidths; w++)
				first[w] = w > st.widx_fixed ? nsc : 0;
		} else {
			choose_widths(ctx, a1, chunk, nsc, ctx->cfg.flags & SAB_FLAG_PER_CHUNK_WIDTHS, b.chunk_host, first);
			if (!b.chunk_host.empty()) {
				rc = ensure(ctx, (void **)&b.chunk_ids, &b.chunk_ids_cap, (size_t)nsc, sizeof(int));
				if (rc)
					return rc;
				CU(cudaMemcpyAsync(b.chunk_ids, b.chunk_host.data(), sizeof(int) * nsc, cudaMemcpyHostToDevice,
						   b.stream));
			}
		}
		const bool listed = st.widx_fixed < 0 && !b.chunk_host.empty();
		WaveVariant var[kNumWidths];
		int max_grid = 0;
		for (int w = 0; w < kNumWidths; w++)
			if (first[w + 1] > first[w]) {
				rc = wave_variant(ctx, method, packed, w, &var[w]);
				if (rc)
					return rc;
				max_grid = std::max(max_grid, var[w].grid);
			}
		int nlaunch = 0;
		for (int w = 0; w < kNumWidths; w++)
			nlaunch += first[w + 1] > first[w];
		const bool serial = ctx->cfg.flags & SAB_FLAG_SERIAL_WIDTHS;
		const size_t bnd_block = (size_t)bnd_rows * 32;
		// concurrent launches need a wrap-around boundary region each (indexed by blockIdx)
		const size_t bnd_launch = bnd_block * max_grid;
		rc = ensure(ctx, (void **)&b.bnd, &b.bnd_cap, bnd_launch * (serial ? 1 : std::max(1, nlaunch)), sizeof(uint2));
		if (rc)
			return rc;
		TileParams p{};
		p.seqs = dseqs;
		p.a_slot0 = st.a0;
		p.a_n = st.an;
		p.a_orig0 = st.a0;
		p.blk = b.blk;
		p.tri0 = tri(st.a0);
		p.blkT = (!triangular && matrix) ? b.blkT : nullptr;
		p.chunk = chunk;
		p.counter = b.counter;
		p.bnd = b.bnd;
		p.bnd_block_stride = (int64_t)bnd_block;
		p.sub25 = ctx->d_sub25;
		p.gapA = ctx->gapA;
		p.gapE = ctx->gapE;
		p.one = 1u;
		// SW a-posteriori test: while every M <= 32767 - max(S) the next diagonal step
		// cannot wrap, so a wrapped pair always shows a running maximum above it first.
		p.ovf_threshold = 32767 - std::max(ctx->max_s, 0);
		p.rerun_list = b.rerun_list;
		p.rerun_count = b.counter + 1;
		p.rerun_cap = rerun_cap;
		CU(cudaMemsetAsync(b.counter, 0, kCounterWords * sizeof(unsigned), b.stream));
		if (S.launches == 0)
			CU(cudaEventRecord(ctx->ev_span[0], b.stream));
		CU(cudaEventRecord(b.ev_start, b.stream));
		if (nlaunch > 1 && !serial)
			CU(cudaEventRecord(b.ev_fork, b.stream));
		// One launch per width in use (widest first).  They touch disjoint subject chunks of the
		// strip buffer, so all but the first go to side streams: the persistent blocks of launch
		// k + 1 move onto the SMs as those of launch k run out of work items.
		int li = 0;
		for (int w = 0; w < kNumWidths; w++) {
			p.nsc = first[w + 1] - first[w];
			if (p.nsc <= 0)
				continue;
			p.chunk_ids = listed ? b.chunk_ids + first[w] : nullptr;
			p.counter = b.counter + 4 + w; // the re-run count (counter[1]) is shared and accumulates
			cudaStream_t s = b.stream;
			if (li > 0 && !serial) {
				if (!b.side[li]) {
					CU(cudaStreamCreateWithFlags(&b.side[li], cudaStreamNonBlocking));
					CU(cudaEventCreateWithFlags(&b.ev_join[li], cudaEventDisableTiming));
				}
				s = b.side[li];
				CU(cudaStreamWaitEvent(s, b.ev_fork, 0));
				p.bnd = b.bnd + bnd_launch * li;
			}
			{
				std::string err;
				if (sab_module_launch(var[w].fn, var[w].grid, var[w].smem, s, &p, &err))
					return fail(ctx, SAB_ERR_CUDA, "%s", err.c_str());
			}
			if (s != b.stream)
				CU(cudaEventRecord(b.ev_join[li], s));
			S.launches++;
			S.width_launches[w]++;
			li++;
		}
		if (!serial)
			for (int k = 1; k < li; k++)
				CU(cudaStreamWaitEvent(b.stream, b.ev_join[k], 0));
		CU(cudaEventRecord(b.ev_done, b.stream));
		S.tiles++;
		if (!packed)
			S.tiles_int32++;
		return SAB_OK;
	};

	auto finish = [&](const Strip &st, StripBuf &b) -> int {
		CU(cudaEventSynchronize(b.ev_done));
		float t_s = 0.f, t_e = 0.f;
		CU(cudaEventElapsedTime(&t_s, ctx->ev_span[0], b.ev_start));
		CU(cudaEventElapsedTime(&t_e, ctx->ev_span[0], b.ev_done));
		const int64_t np = tri(st.a0 + st.an) - tri(st.a0);
		if (method == SAB_SW && st.packed) {
			unsigned cnt = 0;
			CU(cudaMemcpyAsync(&cnt, b.counter + 1, sizeof cnt, cudaMemcpyDeviceToHost, b.stream));
			CU(cudaStreamSynchronize(b.stream));
			// The per-pair re-run (one thread per pair, rows in global scratch) is for the odd
			// overflowing pair.  When the list overflowed, or holds more than ~2 % of the strip
			// (break-even against the int32 instantiation of the strip kernel, which runs at
			// roughly half the packed rate), the whole strip is swept again in int32 instead.
			const bool whole = cnt > rerun_cap || (!test_cap && (int64_t)cnt > std::max<int64_t>(4096, np / 50));
			if (whole) {
				Strip st32 = st;
				st32.packed = false;
				ctx->sw_prefer_int32 = true; // later strips that can overflow start in int32
				int rc = launch(st32, b);
				if (rc)
					return rc;
				CU(cudaEventSynchronize(b.ev_done));
				CU(cudaEventElapsedTime(&t_e, ctx->ev_span[0], b.ev_done));
				S.rerun_strips++;
			} else if (cnt) {
				const int cols = ctx->max_len + 1;
				// one thread per listed pair, at most one block per SM and 512 MB of row scratch
				const size_t per_thread = (size_t)2 * cols * sizeof(int32_t);
				int blocks = (int)std::min<size_t>(((size_t)cnt + kRerunBlock - 1) / kRerunBlock,
								   (size_t)ctx->sm_count);
				blocks = (int)std::max<size_t>(1, std::min<size_t>(blocks, (512u << 20) / (per_thread * kRerunBlock)));
				const int kRerunThreads = blocks * kRerunBlock;
				int rc = ensure(ctx, (void **)&b.rerun_scratch, &b.rerun_scratch_cap,
						(size_t)kRerunThreads * 2 * cols, sizeof(int32_t));
				if (rc)
					return rc;
				sab::RerunParams rp{};
				rp.seqs = dseqs;
				rp.list = b.rerun_list;
				rp.count = b.counter + 1;
				rp.cap = rerun_cap;
				rp.scratch = b.rerun_scratch;
				rp.scratch_cols = cols;
				rp.sub25 = ctx->d_sub25;
				rp.gapA = ctx->gapA;
				rp.gapE = ctx->gapE;
				rp.a_orig0 = st.a0;
				rp.a_n = st.an;
				rp.blk = b.blk;
				rp.tri0 = tri(st.a0);
				rp.blkT = (!triangular && matrix) ? b.blkT : nullptr;
				sab::sab_rerun_sw_kernel<<<blocks, kRerunBlock, 0, b.stream>>>(rp);
				CU(cudaGetLastError());
				CU(cudaEventRecord(b.ev_done, b.stream));
				CU(cudaEventSynchronize(b.ev_done));
				CU(cudaEventElapsedTime(&t_e, ctx->ev_span[0], b.ev_done));
				S.launches++;
				S.rerun_pairs += cnt;
			}
		}
		spans.push_back({ t_s, t_e });
		if (matrix) {
			int rc = drain_strip(ctx, st, b, matrix, triangular);
			if (rc)
				return rc;
		}
		pairs_done += np;
		S.pairs += np;
		if (progress)
			progress(pairs_done, pairs_total, user);
		return SAB_OK;
	};

	// two strips in flight: strip k+1 computes while strip k is re-run / copied back
	Strip inflight[2];
	for (int k = 0, prev = -1;; k++) {
		const int idx = src.next();
		if (idx >= 0) {
			Strip &st = inflight[k & 1];
			st = strips[idx];
			// once a strip of this job had to be swept again in int32, strips that can overflow
			// as well (min(lq, ls) * max(S) above the threshold) skip the packed attempt
			if (method == SAB_SW && st.packed && ctx->sw_prefer_int32 &&
			    (int64_t)std::min(st.lq, st.ls) * std::max(ctx->max_s, 0) > 32767 - std::max(ctx->max_s, 0))
				st.packed = false;
			int rc = launch(st, ctx->buf[k & 1]);
			if (rc)
				return rc;
			mine.push_back(idx);
		}
		if (prev >= 0) {
			int rc = finish(inflight[(k - 1) & 1], ctx->buf[(k - 1) & 1]);
			if (rc)
				return rc;
		}
		prev = idx;
		if (idx < 0)
			break;
	}
	// exact cell count of this call: sum over its strips of sum_{hi} len_hi * (sum_{lo<hi} len_lo)
	{
		std::vector<int32_t> len_orig(num);
		for (int s = 0; s < num; s++)
			len_orig[ctx->slot_orig[s]] = ctx->slot_len[s];
		std::vector<int64_t> pre(num + 1, 0);
		for (int k = 0; k < num; k++)
			pre[k + 1] = pre[k] + len_orig[k];
		int64_t cells = 0;
		for (int idx : mine)
			for (int hi = strips[idx].a0; hi < strips[idx].a0 + strips[idx].an; hi++)
				cells += (int64_t)len_orig[hi] * pre[hi];
		S.cells = cells;
	}
	{ // kernel_ms = length of the union of the kernel intervals (two streams overlap)
		std::sort(spans.begin(), spans.end());
		double uni = 0, busy = 0, cur_s = 0, cur_e = -1;
		for (auto &iv : spans) {
			busy += iv.second - iv.first;
			if (cur_e < 0 || iv.first > cur_e) {
				if (cur_e >= 0)
					uni += cur_e - cur_s;
				cur_s = iv.first;
				cur_e = iv.second;
			} else {
				cur_e = std::max<double>(cur_e, iv.second);
			}
		}
		if (cur_e >= 0)
			uni += cur_e - cur_s;
		S.kernel_ms = uni;
		S.busy_ms = busy;
	}
	if (!mine.empty()) { // everything above has been synchronised; close the device span
		float dev = 0.f;
		CU(cudaEventRecord(ctx->ev_span[1], ctx->copy_stream));
		CU(cudaEventSynchronize(ctx->ev_span[1]));
		CU(cudaEventElapsedTime(&dev, ctx->ev_span[0], ctx->ev_span[1]));
		S.device_ms = dev;
	}
	S.total_ms = now_ms() - t0;
	return SAB_OK;
}

} // namespace

int sab_align_shard(sab_ctx *ctx, int32_t shard, int32_t nshards, int32_t *matrix, int triangular,
		    sab_progress_fn progress, void *user)
{
	if (!ctx)
		return SAB_ERR_ARG;
	if (!ctx->uploaded)
		return fail(ctx, SAB_ERR_STATE, "sab_upload must precede sab_align");
	if (nshards < 1 || shard < 0 || shard >= nshards)
		return fail(ctx, SAB_ERR_ARG, "bad shard %d of %d", shard, nshards);
	std::vector<Strip> strips;
	std::vector<int> owner, mine;
	plan_strips(ctx, nshards, strips, owner);
	for (int idx = 0; idx < (int)strips.size(); idx++)
		if (owner[idx] == shard)
			mine.push_back(idx);
	issue_order(ctx, strips, mine);
	int64_t pairs_total = 0;
	for (int idx : mine)
		pairs_total += tri(strips[idx].a0 + strips[idx].an) - tri(strips[idx].a0);
	std::atomic<int> cursor{ 0 };
	StripSource src{ &mine, &cursor };
	return align_strips(ctx, strips, src, pairs_total, matrix, triangular, progress, user);
}

int sab_align(sab_ctx *ctx, int32_t *matrix, int triangular, sab_progress_fn progress, void *user)
{
	return sab_align_shard(ctx, 0, 1, matrix, triangular, progress, user);
}

int sab_plan_strips(const sab_meta *meta, int32_t num, int32_t panel, int32_t nshards,
		    int32_t max_strips, int32_t *row0, int32_t *rows, int32_t *owner_out)
{
	if (!meta || num < 2 || nshards < 1 || panel < 0 || panel % kPanelQuantum)
		return SAB_ERR_ARG;
	std::vector<int64_t> P((size_t)num + 1, 0);
	for (int32_t k = 0; k < num; k++)
		P[k + 1] = P[k] + meta[k].len;
	std::vector<int32_t> start;
	panel_bounds(P, num, panel, nshards, false, false, start);
	const int npanels = (int)start.size() - 1;
	if (npanels > max_strips)
		return npanels;
	std::vector<double> cost(npanels);
	for (int p = 0; p < npanels; p++) {
		const int a0 = start[p], a1 = start[p + 1];
		const double lq = (double)(P[a1] - P[a0]);
		cost[p] = lq * ((double)P[a0] + 0.5 * lq);
		if (row0)
			row0[p] = a0;
		if (rows)
			rows[p] = a1 - a0;
	}
	std::vector<int> owner;
	lpt_assign(cost, nshards, owner);
	if (owner_out)
		for (int p = 0; p < npanels; p++)
			owner_out[p] = owner[p];
	return npanels;
}

int sab_strip_count(const sab_ctx *ctx)
{
	if (!ctx || !ctx->uploaded)
		return SAB_ERR_STATE;
	return (int)ctx->panel_start.size() - 1;
}

int sab_strip_info(const sab_ctx *ctx, int32_t strip, int32_t nshards, int32_t *row0, int32_t *rows,
		   int32_t *owner_out)
{
	if (!ctx || !ctx->uploaded)
		return SAB_ERR_STATE;
	if (nshards < 1 || strip < 0 || strip >= sab_strip_count(ctx))
		return SAB_ERR_ARG;
	std::vector<Strip> strips;
	std::vector<int> owner;
	plan_strips(ctx, nshards, strips, owner);
	if (row0)
		*row0 = strips[strip].a0;
	if (rows)
		*rows = strips[strip].an;
	if (owner_out)
		*owner_out = owner[strip];
	return SAB_OK;
}

int sab_get_stats(const sab_ctx *ctx, sab_stats *out)
{
	if (!ctx || !out)
		return SAB_ERR_ARG;
	*out = ctx->stats;
	return SAB_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------
// All GPUs of the box from ONE host process: one context + one host thread per device,
// every device holds a replica of the sequence set and pulls row strips from a cursor shared
// by the devices (costliest first), writing them straight into the caller's matrix (disjoint
// rows -> no locking, no collective).  This is what the reference-signature shim
// (integration/seqalign_shim.c) drives.
//
// Contexts are created lazily and in parallel, and only as many as the job can keep busy: a CUDA
// context costs a few hundred ms of host time, a C2-sized job (4.5e12 cells) takes 0.5 s on one
// device (round 1: the drop-in CLI took 1.7 s with one visible device and 5.2 s with eight).
// ---------------------------------------------------------------------------------------
#include <mutex>

struct sab_multi {
	sab_config cfg{};
	std::vector<int> ordinals; // usable CUDA ordinals this object may drive
	int want = 0; // devices the caller asked for (0 = size from the job)
	std::vector<sab_ctx *> ctx; // created so far, ctx[k] drives ordinals[k]
	int active = 0; // contexts holding the current sequence set
	std::string err;
	sab_stats stats{};
	// progress aggregation
	std::mutex mu;
	std::vector<int64_t> done;
	int64_t total = 0;
	sab_progress_fn progress = nullptr;
	void *user = nullptr;
};

namespace {

struct MultiSlot {
	sab_multi *m;
	int k;
};

void multi_progress(int64_t done, int64_t, void *user)
{
	MultiSlot *s = static_cast<MultiSlot *>(user);
	sab_multi *m = s->m;
	std::lock_guard<std::mutex> lock(m->mu);
	m->done[s->k] = done;
	int64_t sum = 0;
	for (int64_t d : m->done)
		sum += d;
	if (m->progress)
		m->progress(sum, m->total, m->user);
}

template <typename F> int multi_run(sab_multi *m, int n, F fn)
{
	std::vector<int> rc(n, SAB_OK);
	if (n == 1) {
		rc[0] = fn(0);
	} else {
		std::vector<std::thread> th;
		for (int k = 0; k < n; k++)
			th.emplace_back([&, k]() { rc[k] = fn(k); });
		for (auto &t : th)
			t.join();
	}
	for (int k = 0; k < n; k++)
		if (rc[k] != SAB_OK) {
			m->err = "device " + std::to_string(m->ordinals[k]) + ": " +
				 (k < (int)m->ctx.size() && m->ctx[k] ? sab_last_error(m->ctx[k]) : "no context");
			return rc[k];
		}
	return SAB_OK;
}

// Makes sure contexts [0, n) exist; the missing ones are created concurrently.
int multi_ensure(sab_multi *m, int n)
{
	const int have = (int)m->ctx.size();
	if (n <= have)
		return SAB_OK;
	m->ctx.resize(n, nullptr);
	std::vector<int> rc(n, SAB_OK);
	std::vector<std::thread> th;
	auto make = [&](int k) {
		sab_config c = m->cfg;
		c.device = m->ordinals[k];
		c.shards_hint = std::max(m->want, (int)m->ordinals.size());
		c.flags |= SAB_FLAG_INTERNAL_DYNAMIC; // strips are pulled from a shared cursor: uniform panels
		if (c.host_threads > 0) // the caller's thread budget is shared by the devices
			c.host_threads = std::max(1, c.host_threads / std::max(1, n));
		rc[k] = sab_create(&m->ctx[k], &c);
	};
	if (n - have == 1) {
		make(have);
	} else {
		for (int k = have; k < n; k++)
			th.emplace_back(make, k);
		for (auto &t : th)
			t.join();
	}
	for (int k = have; k < n; k++)
		if (rc[k] != SAB_OK) {
			m->err = "device " + std::to_string(m->ordinals[k]) + ": " + sab_last_error(m->ctx[k]);
			const int code = rc[k];
			for (int j = have; j < n; j++)
				sab_destroy(m->ctx[j]);
			m->ctx.resize(have);
			return code;
		}
	m->done.assign(n, 0);
	return SAB_OK;
}

// Devices worth starting for `seconds` of single-device work: each one should get at least
// ~0.4 s of it (what creating its context and replicating the sequence set costs the host).
int devices_for(const sab_multi *m, double seconds)
{
	const int cap = (int)m->ordinals.size();
	if (m->want > 0)
		return std::min(m->want, cap);
	return std::max(1, std::min(cap, (int)(seconds / 0.4)));
}

} // namespace

extern "C" {

int sab_multi_create(sab_multi **out, const sab_config *cfg, int32_t ndev)
{
	if (!out || !cfg)
		return SAB_ERR_ARG;
	*out = nullptr;
	sab_multi *m = new (std::nothrow) sab_multi;
	if (!m)
		return SAB_ERR_NOMEM;
	*out = m;
	m->cfg = *cfg;
	m->ordinals = usable_devices();
	const int have = (int)m->ordinals.size();
	if (have == 0 || ndev > have) {
		m->err = have ? "more devices requested than present" :
				"no CUDA device (this library has no CPU fallback)";
		return SAB_ERR_NODEVICE;
	}
	m->want = ndev > 0 ? ndev : 0;
	// validate the configuration now (sab_create's argument checks), on the first device
	return multi_ensure(m, 1);
}

void sab_multi_destroy(sab_multi *m)
{
	if (!m)
		return;
	std::vector<std::thread> th; // context teardown is as slow as creation: in parallel too
	for (sab_ctx *c : m->ctx)
		th.emplace_back([c]() { sab_destroy(c); });
	for (auto &t : th)
		t.join();
	delete m;
}

const char *sab_multi_last_error(const sab_multi *m)
{
	return m ? m->err.c_str() : "null context";
}

int sab_multi_device_count(const sab_multi *m)
{
	if (!m)
		return 0;
	return m->active ? m->active : (m->want ? m->want : (int)m->ordinals.size());
}

int sab_multi_upload(sab_multi *m, const uint8_t *seqs, const sab_meta *meta, int32_t num, int32_t max)
{
	if (!m || m->ctx.empty() || !meta || num < 2)
		return SAB_ERR_ARG;
	// size the device set from the job: cells / ~6 TCUPS per device
	double sum = 0, sq = 0;
	for (int32_t k = 0; k < num; k++) {
		sum += meta[k].len;
		sq += (double)meta[k].len * meta[k].len;
	}
	const int n = devices_for(m, (sum * sum - sq) / 2 / 6e12);
	int rc = multi_ensure(m, n);
	if (rc)
		return rc;
	m->active = 0;
	// packed once (the auto panel sized for the devices that take part in THIS job), copied to every device
	const double t0 = now_ms();
	m->ctx[0]->cfg.shards_hint = n;
	PackedSet ps;
	rc = pack_set(m->ctx[0], seqs, meta, num, max, ps);
	if (rc) {
		m->err = sab_last_error(m->ctx[0]);
		return rc;
	}
	rc = multi_run(m, n, [&](int k) {
		m->ctx[k]->cfg.shards_hint = n;
		const int r = install_set(m->ctx[k], ps, true);
		m->ctx[k]->stats.upload_ms = now_ms() - t0;
		return r;
	});
	if (rc == SAB_OK)
		m->active = n;
	return rc;
}

int sab_multi_align(sab_multi *m, int32_t *matrix, int triangular, sab_progress_fn progress, void *user)
{
	if (!m || m->ctx.empty())
		return SAB_ERR_ARG;
	const int n = m->active;
	if (n < 1) {
		m->err = "sab_multi_upload must precede sab_multi_align";
		return SAB_ERR_STATE;
	}
	m->progress = progress;
	m->user = user;
	m->done.assign(n, 0);
	m->total = tri(m->ctx[0]->num);
	std::vector<MultiSlot> slots(n);
	for (int k = 0; k < n; k++)
		slots[k] = { m, k };
	// one plan (every context holds the same set and panel), one cursor: dynamic balancing
	std::vector<Strip> strips;
	std::vector<int> owner, order;
	plan_strips(m->ctx[0], 1, strips, owner);
	order.resize(strips.size());
	std::iota(order.begin(), order.end(), 0);
	issue_order(m->ctx[0], strips, order);
	std::atomic<int> cursor{ 0 };
	const double t0 = now_ms();
	const int rc = multi_run(m, n, [&](int k) {
		StripSource src{ &order, &cursor };
		return align_strips(m->ctx[k], strips, src, m->total, matrix, triangular, multi_progress,
				    &slots[k]);
	});
	if (rc != SAB_OK)
		return rc;
	sab_stats S{};
	for (int k = 0; k < n; k++) { // counts add up; times are the slowest device's
		const sab_stats &s = m->ctx[k]->stats;
		S.kernel_ms = std::max(S.kernel_ms, s.kernel_ms);
		S.busy_ms = std::max(S.busy_ms, s.busy_ms);
		S.device_ms = std::max(S.device_ms, s.device_ms);
		S.upload_ms = std::max(S.upload_ms, s.upload_ms);
		S.pairs += s.pairs;
		S.cells += s.cells;
		S.rerun_pairs += s.rerun_pairs;
		S.rerun_strips += s.rerun_strips;
		for (int w = 0; w < 16; w++)
			S.width_launches[w] += s.width_launches[w];
		S.launches += s.launches;
		S.tiles += s.tiles;
		S.tiles_int32 += s.tiles_int32;
		S.h2d_bytes += s.h2d_bytes;
		S.d2h_bytes += s.d2h_bytes;
	}
	S.filter_ms = m->stats.filter_ms;
	S.filter_matches = m->stats.filter_matches;
	S.total_ms = now_ms() - t0;
	m->stats = S;
	return SAB_OK;
}

int sab_multi_filter(sab_multi *m, const uint8_t *seqs, const sab_meta *meta, int32_t num,
		     float threshold, uint8_t *lost)
{
	if (!m || m->ctx.empty() || !seqs || !meta || !lost || num < 1)
		return SAB_ERR_ARG;
	const double t0 = now_ms();
	// ~1.5e11 prefix compares per second and device (the kernel stops a pair as soon as the
	// threshold is out of reach; C4: 5e9 pairs)
	const int n = devices_for(m, (double)num * (num - 1) / 2 / 1.5e11);
	int rc = multi_ensure(m, n);
	if (rc)
		return rc;
	std::vector<std::vector<uint64_t>> part(n);
	std::vector<int> overflow(n, 0);
	rc = multi_run(m, n, [&](int k) {
		return sab_internal_filter_matches(m->ctx[k], seqs, meta, num, threshold, k, n, part[k],
						   &overflow[k]);
	});
	if (rc != SAB_OK)
		return rc;
	int nlost;
	if (std::any_of(overflow.begin(), overflow.end(), [](int o) { return o != 0; })) {
		// too many matching pairs to list (a threshold most pairs pass): the exact
		// panel-sequential path on one device
		nlost = sab_filter(m->ctx[0], seqs, meta, num, threshold, lost);
		if (nlost < 0) {
			m->err = sab_last_error(m->ctx[0]);
			return nlost;
		}
	} else {
		std::vector<uint64_t> all;
		for (auto &p : part)
			all.insert(all.end(), p.begin(), p.end());
		nlost = sab_filter_resolve(num, all.data(), (int64_t)all.size(), lost);
		m->stats.filter_matches = (int64_t)all.size();
	}
	m->stats.filter_ms = now_ms() - t0;
	return nlost;
}

int sab_multi_get_stats(const sab_multi *m, sab_stats *out)
{
	if (!m || !out)
		return SAB_ERR_ARG;
	*out = m->stats;
	return SAB_OK;
}

} // extern "C"
