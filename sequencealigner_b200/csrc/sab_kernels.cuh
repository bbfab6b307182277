// sab_kernels.cuh -- shared device definitions of the sm_100a score-only all-vs-all DP fill.
//
// Replaces (does not port) the reference's one-thread-per-pair int32 kernels
// kernel_nw / kernel_ga / kernel_sw (reference: src/bio/kernels.cu:32-221), whose DP
// rows live in 8-20 KB of per-thread local memory and whose every cell does
// uncoalesced global + constant loads.  The kernel itself is in sab_wave.cuh; this file
// holds what it shares with the host code and the int32 re-run kernel:
//
//  * DeviceSeqs / TileParams: the device-resident sequence store and the launch record
//    of one row strip.
//  * VOps: two int16 lanes per 32-bit register (Blackwell DPX: VIADDMNMX.U16x2,
//    VIMNMX3.U16x2) or one int32 lane.  Measured on B200 (tools/pipe_peak.cu): the DPX
//    ops, IADD3 and the packed VIADD.16x2 all issue on the 16-lane alu pipe, IMAD on the
//    equally wide fma pipe.  The int16 lanes are therefore kept BIASED (value + 0x8000,
//    compared unsigned): a plain 32-bit add of an arithmetic-encoded delta
//    (d1 * 65536 + d0) then updates both lanes without any carry between them as long as
//    the low lane stays inside [0, 65535], which the range test guarantees, and those adds
//    go to the fma pipe as IMAD (x * 1 + y with the 1 in a uniform register).
//  * int16 range: decided a priori per row strip for the global methods; SW detects
//    overflow a posteriori from its own running maximum and queues the pair for the
//    int32 re-run kernel below.  The int32 instantiation of the main kernel (one subject
//    per pass) serves strips that fail the a-priori test.
//
// Recurrences and borders follow the reference CPU code exactly:
//   NW  src/bio/method/nw.c:11-42,  GA  src/bio/method/ga.c:23-67,  SW  src/bio/method/sw.c:15-61.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace sab {

constexpr int kStrip = 32; // widest register strip (DP columns per warp); subjects are padded to it at most
constexpr int kLetters = 25; // 24 residues + 1 pad letter
constexpr int kProfStride = 25; // words per profile column (coprime with 32 banks)

enum { M_NW = 0, M_GA = 1, M_SW = 2 };

constexpr int kWaveWarps = 4; // warps of a block = consecutive strips of the strip pipeline (sab_wave.cuh)
constexpr int kWaveThreads = kWaveWarps * 32;

// what the profile entries are shifted by (host: sub25 = S - shift; device: the skewed cell of sab_wave.cuh)
__host__ __device__ inline int profile_shift(int method, int gapA, int gapE)
{
	return method == M_NW ? 2 * gapA : method == M_GA ? 2 * gapE : gapA;
}

// Device-resident sequence store (built by sab_upload).
struct DeviceSeqs {
	const int32_t *slot_len; // [nslots]   length of the sequence in slot s
	const int32_t *slot_orig; // [nslots]   original index of the sequence in slot s
	const int64_t *col_off; // [nslots]   word offset of the slot's letters in col_words
	const uint32_t *col_words; // subject letters as a bit stream per slot, col_bits bits per letter
	const int64_t *grp_off; // [ngroups]  row4 offset of the 32-slot group in rowpack
	const uint32_t *rowpack; // [(grp_off[g] + r/4) * 32 + lane] = 4 letters * 4 (byte offsets)
	int col_bits; // 2: the set uses <= 4 letters of the alphabet (ACGT reads), codes via col_dec; 5: letter index
	uint32_t col_dec; // 2-bit code -> letter index, one byte per code
};

// Letter index (0..23) of residue `pos` of the sequence in `slot`.
__device__ __forceinline__ int subject_letter(const DeviceSeqs &q, const int slot, const int pos)
{
	const uint32_t *w = q.col_words + q.col_off[slot];
	const int b = pos * q.col_bits;
	const int sh = b & 31;
	uint32_t v = w[b >> 5] >> sh;
	if (sh + q.col_bits > 32) // a 5-bit letter may straddle two words (of the same slot)
		v |= w[(b >> 5) + 1] << (32 - sh);
	if (q.col_bits == 2)
		return (int)((q.col_dec >> (8 * (v & 3u))) & 0xffu);
	return (int)(v & 31u);
}

// One launch = one ROW STRIP of the score matrix: the queries are the slots of panel
// `a` (original indices [a_orig0, a_orig0 + a_n)), the subjects every slot below them.
// Scores land in the strip buffer in the reference's packed-triangular row layout
// (src/io/output.c:79-83): blk[tri(hi) - tri(a_orig0) + lo], tri(j) = j(j-1)/2, hi > lo
// original indices -- so the strip is one contiguous run of the final matrix.
struct TileParams {
	DeviceSeqs seqs;
	int a_slot0, a_n; // query slots [a_slot0, a_slot0 + a_n); subjects: slots [0, a_slot0 + a_n)
	int a_orig0; // first original index of the strip (== a_slot0: panels are index ranges)
	int32_t *blk; // strip buffer, packed-triangular rows
	int64_t tri0; // tri(a_orig0)
	int32_t *blkT; // optional mirror for the full layout: blkT[lo * a_n + (hi - a_orig0)]
	int nsc, chunk; // work items = query groups of 32 x nsc subject chunks of `chunk` slots
	const int *chunk_ids; // the nsc chunks this launch covers (nullptr: chunks 0..nsc-1); a strip is
			      // launched once per register-strip width, each launch taking the chunks
			      // whose subjects pad least at that width
	unsigned *counter; // dynamic work counter of this launch
	uint2 *bnd; // wrap-around boundary rows, one buffer per block
	int64_t bnd_block_stride; // uint2 elements per block
	const int32_t *sub25; // [25*25] S[a][q] - shift(method), pad row/col = -shift (sab_wave.cuh)
	int gapA; // NW: gap penalty, GA/SW: gap open (<= 0)
	int gapE; // GA/SW: gap extend (<= 0)
	uint32_t one; // == 1, kept opaque so the packed adds issue as IMAD on the fma pipe
	int ovf_threshold; // SW int16: a score above this may have wrapped -> re-run
	unsigned long long *rerun_list; // (query slot << 32 | subject slot)
	unsigned *rerun_count;
	unsigned rerun_cap;
};

// Register-strip widths the strip kernel is built for: one device image per width (sab_wave_inst.cu,
// compiled to a cubin and embedded by build.py; loaded and specialised by sab_module.cu).
#define SAB_WIDTHS(X) X(32) X(30) X(28) X(26) X(24) X(22) X(20) X(18) X(16)
// resident blocks per SM the register budget allows: 2K registers hold the strip itself
constexpr int wave_minb(int k)
{
	return k <= 20 ? 5 : 4;
}
// placeholder of the per-lane gap constant in the built images (sab_module.cu overwrites it)
constexpr uint32_t kLgPlaceholder = 0x5ab75ab7u;

// ---------------------------------------------------------------------------------------
// Vector ops: two int16 lanes per register (Blackwell DPX) or one int32.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t imad_add(uint32_t a, uint32_t one, uint32_t b)
{
	uint32_t d; // a * 1 + b on the fma pipe; `one` lives in a uniform register
	asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
	return d;
}

template <bool PACKED> struct VOps;

// Two biased int16 lanes (stored value = score + 0x8000, compared unsigned).
template <> struct VOps<true> {
	static constexpr int kBias = 0x8000;
	// both lanes += the same / a per-lane signed delta, as ONE 32-bit add
	static __device__ __forceinline__ uint32_t delta(int v) { return (uint32_t)(v * 0x10001); }
	static __device__ __forceinline__ uint32_t delta2(int lo, int hi)
	{
		return (uint32_t)(hi * 65536 + lo);
	}
	static __device__ __forceinline__ uint32_t addw(uint32_t a, uint32_t d, uint32_t one)
	{
		return imad_add(a, one, d);
	}
	// two per-lane signed addends for the DPX add-max ops (each wraps inside its own lane)
	static __device__ __forceinline__ uint32_t lane2(int lo, int hi)
	{
		return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16);
	}
	// per-lane addend for the DPX add-max ops (wraps inside each lane)
	static __device__ __forceinline__ uint32_t lane(int v)
	{
		return ((uint32_t)v & 0xffffu) * 0x10001u;
	}
	static __device__ __forceinline__ uint32_t addmax(uint32_t a, uint32_t b, uint32_t c)
	{
		return __viaddmax_u16x2(a, b, c);
	}
	static __device__ __forceinline__ uint32_t max3(uint32_t a, uint32_t b, uint32_t c)
	{
		return __vimax3_u16x2(a, b, c);
	}
	static __device__ __forceinline__ uint32_t bcast(int v) { return lane(v + kBias); }
	// -inf stand-in: NEG + ext == biased 0 exactly, so the first use can never wrap and
	// always loses the max against a real (range-checked) value.
	static __device__ __forceinline__ uint32_t neg_inf(int ext) { return lane(-ext); }
	static __device__ __forceinline__ int lo(uint32_t v) { return (int)(v & 0xffffu) - kBias; }
	static __device__ __forceinline__ int hi(uint32_t v) { return (int)(v >> 16) - kBias; }
};

// One int32 lane, the reference's own arithmetic (SCORE_MIN sentinel, src/bio/align.h:19).
template <> struct VOps<false> {
	static __device__ __forceinline__ uint32_t delta(int v) { return (uint32_t)v; }
	static __device__ __forceinline__ uint32_t delta2(int lo, int) { return (uint32_t)lo; }
	static __device__ __forceinline__ uint32_t lane2(int lo, int) { return (uint32_t)lo; }
	static __device__ __forceinline__ uint32_t addw(uint32_t a, uint32_t d, uint32_t one)
	{
		return imad_add(a, one, d);
	}
	static __device__ __forceinline__ uint32_t lane(int v) { return (uint32_t)v; }
	static __device__ __forceinline__ uint32_t addmax(uint32_t a, uint32_t b, uint32_t c)
	{
		return (uint32_t)__viaddmax_s32((int)a, (int)b, (int)c);
	}
	static __device__ __forceinline__ uint32_t max3(uint32_t a, uint32_t b, uint32_t c)
	{
		return (uint32_t)__vimax3_s32((int)a, (int)b, (int)c);
	}
	static __device__ __forceinline__ uint32_t bcast(int v) { return (uint32_t)v; }
	static __device__ __forceinline__ uint32_t neg_inf(int) { return (uint32_t)(INT32_MIN / 2); }
	static __device__ __forceinline__ int lo(uint32_t v) { return (int)v; }
	static __device__ __forceinline__ int hi(uint32_t) { return 0; }
};

// ---------------------------------------------------------------------------------------
// int32 re-run of individual pairs whose int16 SW score may have wrapped.  One thread per
// queued pair, two rolling rows (M and Y) in global scratch, interleaved across threads
// (element c of thread t at [c * threads + t]: a warp's accesses coalesce).  Rare path (needs a
// score > 32767 - max(S)); lists longer than ~2 % of a strip go to the int32 instantiation of
// the strip kernel instead (sab_capi.cu: finish).
// ---------------------------------------------------------------------------------------
struct RerunParams {
	DeviceSeqs seqs;
	const unsigned long long *list;
	const unsigned *count;
	unsigned cap;
	int32_t *scratch; // [2][scratch_cols][threads]
	int scratch_cols;
	const int32_t *sub25; // as TileParams (offset by -gapA)
	int gapA, gapE;
	int a_orig0, a_n;
	int32_t *blk;
	int64_t tri0;
	int32_t *blkT;
};

#ifdef SAB_WITH_RERUN_KERNEL // defined by the one translation unit that launches it (sab_capi.cu)
__global__ void sab_rerun_sw_kernel(const RerunParams p)
{
	const unsigned n = min(*p.count, p.cap);
	const unsigned nthreads = gridDim.x * blockDim.x;
	const unsigned me = blockIdx.x * blockDim.x + threadIdx.x;
	int32_t *m = p.scratch + me;
	int32_t *y = m + (int64_t)p.scratch_cols * nthreads;
	const int64_t st = nthreads; // stride between consecutive columns of one thread's row
	const int o = p.gapA, e = p.gapE;
	const int NEG = INT32_MIN / 2;
	for (unsigned k = me; k < n; k += nthreads) {
		const unsigned long long ent = p.list[k];
		const int t = (int)(ent >> 32), s = (int)(ent & 0xffffffffu);
		const int Lq = p.seqs.slot_len[t], Ls = p.seqs.slot_len[s];
		for (int c = 0; c <= Ls; c++) {
			m[c * st] = 0;
			y[c * st] = NEG;
		}
		int best = 0;
		for (int r = 0; r < Lq; r++) {
			const int32_t *srow = p.sub25 + subject_letter(p.seqs, t, r) * kLetters;
			int diag = m[0], left = 0, x = NEG;
			for (int c = 1; c <= Ls; c++) {
				const int sub = srow[subject_letter(p.seqs, s, c - 1)] + o; // sub25 carries -gapA
				const int d = diag + sub;
				const int up = m[c * st];
				x = max(left + o, x + e);
				const int yy = max(up + o, y[c * st] + e);
				int v = max(d, 0);
				v = max(v, x);
				v = max(v, yy);
				diag = up;
				m[c * st] = v;
				y[c * st] = yy;
				left = v;
				best = max(best, v);
			}
		}
		const int qo = p.seqs.slot_orig[t], so = p.seqs.slot_orig[s];
		const int hi = max(qo, so), lo = min(qo, so);
		p.blk[(int64_t)hi * (hi - 1) / 2 - p.tri0 + lo] = best;
		if (p.blkT)
			p.blkT[(int64_t)lo * p.a_n + (hi - p.a_orig0)] = best;
	}
}
#endif // SAB_WITH_RERUN_KERNEL

} // namespace sab
