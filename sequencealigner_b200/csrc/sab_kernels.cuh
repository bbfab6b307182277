// sab_kernels.cuh -- hand-written sm_100a kernels for the score-only all-vs-all DP fill.
//
// Replaces (does not port) the reference's one-thread-per-pair int32 kernels
// kernel_nw / kernel_ga / kernel_sw (reference: src/bio/kernels.cu:32-221), whose DP
// rows live in 8-20 KB of per-thread local memory and whose every cell does
// uncoalesced global + constant loads.  Design (DESIGN.md sections 3-5):
//
//  * One thread owns one QUERY sequence (DP rows, streamed letter by letter) and the
//    whole block shares one SUBJECT PAIR (DP columns): the two subjects are packed into
//    the low/high int16 halves of every 32-bit register, so each Blackwell DPX
//    instruction (VIADDMNMX.S16x2 / VIMNMX3.S16x2[.RELU] / VIADD.16x2) updates two cells.
//  * The subject pair is expanded once per block into a shared-memory PAIR PROFILE
//    prof[column][query letter] = (S[a0_col][q], S[a1_col][q]) - gap offset.  All 32
//    lanes of a warp read the same profile column with their own query letter: 25
//    consecutive words -> never a bank conflict, address = per-row base + immediate.
//  * A thread keeps a strip of kStrip=32 columns (Mo = M+open, Y) in registers and
//    sweeps all query rows; strips of one subject pair hand their last column
//    (M+open, X) to the next strip through a per-thread, warp-interleaved (coalesced)
//    boundary buffer in global memory that stays L2 resident.
//  * Five integer instructions per packed cell pair for Gotoh/SW, three for NW:
//        Y  = viaddmax(Y, ext, Mo_up)         X = viaddmax(X, ext, Mo_left)
//        d  = Mo_diag + (s - open)            M = vimax3(d, X, Y)
//        Mo = M + open                        (+ 1/2 vimax3 per cell for SW's running max)
//    Measured on B200 (tools/pipe_peak.cu): the DPX ops, IADD3 and the packed VIADD.16x2
//    all issue on the 16-lane alu pipe, IMAD on the equally wide fma pipe.  The int16
//    lanes are therefore kept BIASED (value + 0x8000, unsigned): a plain 32-bit add of
//    an arithmetic-encoded delta (d1 * 65536 + d0) then updates both lanes without any
//    carry between them as long as the low lane stays inside [0, 65535], which the range
//    test guarantees.  The two plain adds go to the fma pipe as IMAD (x * 1 + y with the
//    1 in a uniform register), leaving three DPX ops per packed cell on the alu pipe.
//  * int16 range: decided a priori per panel tile for the global methods; SW detects
//    overflow a posteriori from its own running maximum and queues the pair for the
//    int32 re-run kernel.  The int32 instantiation of the same kernel (one subject per
//    pass) serves tiles that fail the a-priori test.
//
// Recurrences and borders follow the reference CPU code exactly:
//   NW  src/bio/method/nw.c:11-42,  GA  src/bio/method/ga.c:23-67,  SW  src/bio/method/sw.c:15-61.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace sab {

constexpr int kThreads = 128; // queries (threads) per block
constexpr int kStrip = 32; // DP columns held in registers per strip
constexpr int kWindow = 128; // profile columns staged per unit (4 strips)
constexpr int kLetters = 25; // 24 residues + 1 pad letter
constexpr int kProfStride = 25; // words per profile column (coprime with 32 banks)
constexpr int kSmemBytes = (2 * kWindow * kProfStride + kLetters * kLetters) * 4 + 16;

enum { M_NW = 0, M_GA = 1, M_SW = 2 };

// Device-resident sequence store (built by sab_upload).
struct DeviceSeqs {
	const int32_t *slot_len; // [nslots]   length of the sequence in slot s
	const int32_t *slot_orig; // [nslots]   original index of the sequence in slot s
	const int64_t *col_off; // [nslots]   byte offset of the slot's letters in col_bytes
	const uint8_t *col_bytes; // letter indices 0..23, one byte per residue
	const int64_t *grp_off; // [ngroups]  row4 offset of the 32-slot group in rowpack
	const uint32_t *rowpack; // [(grp_off[g] + r/4) * 32 + lane] = 4 letters * 4 (byte offsets)
};

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
	int nqb, nsc, chunk; // work items = nqb query blocks x nsc subject chunks of `chunk` slots
	unsigned *counter; // dynamic work counter of this launch
	uint2 *bnd; // strip boundary buffers
	int64_t bnd_block_stride; // uint2 elements per block
	int bnd_rows; // rows per warp buffer
	const int32_t *sub25; // [25*25] S[a][q] - gapA, pad row/col = -gapA
	int gapA; // NW: gap penalty, GA/SW: gap open (<= 0)
	int gapE; // GA/SW: gap extend (<= 0)
	uint32_t one; // == 1, kept opaque so the packed adds issue as IMAD on the fma pipe
	int ovf_threshold; // SW int16: a score above this may have wrapped -> re-run
	unsigned long long *rerun_list; // (query slot << 32 | subject slot)
	unsigned *rerun_count;
	unsigned rerun_cap;
};

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
// One register strip: columns [c0, c0+kStrip) of the subject (pair) against all Lq rows
// of this thread's query.
//   prof_s : shared profile, column c0 of the current window (stride kProfStride words)
//   rowp   : this lane's packed query letters (one word = 4 rows, stride 32 words)
//   bnd    : this lane's boundary column (stride 32 uint2 per row), read+written in place
// ---------------------------------------------------------------------------------------
template <int METHOD, bool PACKED>
__device__ __forceinline__ void run_strip(const uint32_t *__restrict__ prof_s,
					  const uint32_t *__restrict__ rowp, uint2 *__restrict__ bnd,
					  const int Lq, const int c0, const bool first, const bool last,
					  const int gA, const int gE, const uint32_t one, uint32_t &best,
					  uint32_t &res, const int cap0, const int cap1)
{
	using O = VOps<PACKED>;
	constexpr int K = kStrip;
	const uint32_t dA = O::delta(gA); // 32-bit-add form of "+ gapA" on every lane
	const uint32_t lE = O::lane(gE); // per-lane form of "+ gapE" for the DPX add-max
	const uint32_t NEG = O::neg_inf(gE);
	const uint32_t ZERO = O::bcast(0); // SW floor

	// Row 0 of the DP (borders): Mo[c] = M[0][C] + gapA with C = c0 + c + 1.
	//   NW  M[0][C] = C*g               (nw.c:13-14)
	//   GA  M[0][C] = open + (C-1)*ext  (ga.c:26-30)      Y[0][C] = -inf
	//   SW  M[0][C] = 0                 (sw.c:21-24)      Y[0][C] = -inf
	// SW keeps Y floored at 0 (Yf = max(Y, 0)): M = max(0, d, X, Y) = max(d, X, Yf) and
	// Yf' = max(Yf + ext, Mo_up, 0) is exact for the floored value, so the zero floor costs
	// no extra alu-pipe instruction.
	uint32_t Mo[K], Y[K];
	{
		int m0, st;
		if (METHOD == M_NW) {
			m0 = (c0 + 2) * gA;
			st = gA;
		} else if (METHOD == M_GA) {
			m0 = 2 * gA + c0 * gE;
			st = gE;
		} else {
			m0 = gA;
			st = 0;
		}
		const uint32_t dst = O::delta(st);
		Mo[0] = O::bcast(m0);
		Y[0] = (METHOD == M_SW) ? ZERO : NEG;
#pragma unroll
		for (int c = 1; c < K; c++) {
			Mo[c] = Mo[c - 1] + dst;
			Y[c] = Y[0];
		}
	}
	// diag for row 1 = M[0][c0] + gapA; left border generator for the first strip:
	// bl(R) = M[R][0] + gapA, R = 1.. (nw.c:16-17, ga.c:32-38, sw.c:26-30)
	uint32_t diag, bl, dstep;
	{
		int d0, b0, st;
		if (METHOD == M_NW) {
			d0 = (c0 + 1) * gA;
			b0 = 2 * gA;
			st = gA;
		} else if (METHOD == M_GA) {
			d0 = c0 ? 2 * gA + (c0 - 1) * gE : gA;
			b0 = 2 * gA;
			st = gE;
		} else {
			d0 = gA;
			b0 = gA;
			st = 0;
		}
		diag = O::bcast(d0);
		bl = O::bcast(b0);
		dstep = O::delta(st);
	}

	auto step = [&](const uint32_t qoff, const uint2 bin, const int r) {
		uint32_t ml, x;
		if (first) {
			ml = bl;
			x = NEG;
			bl += dstep;
		} else {
			ml = bin.x;
			x = bin.y;
		}
		const uint32_t *pq = reinterpret_cast<const uint32_t *>(
			reinterpret_cast<const char *>(prof_s) + qoff);
		uint32_t dg = diag;
		diag = ml;
		uint32_t mprev = 0;
#pragma unroll
		for (int c = 0; c < K; c++) {
			const uint32_t s = pq[c * kProfStride];
			const uint32_t t = Mo[c];
			const uint32_t d = O::addw(dg, s, one); // fma pipe
			uint32_t m;
			if (METHOD == M_NW) {
				m = O::max3(d, ml, t);
			} else {
				uint32_t y;
				if (METHOD == M_SW) {
					if (PACKED) {
						const uint32_t ye = O::addw(Y[c], O::delta(gE), one); // fma pipe
						y = O::max3(ye, t, ZERO);
					} else {
						y = (uint32_t)__viaddmax_s32_relu((int)Y[c], gE, (int)t);
					}
				} else {
					y = O::addmax(Y[c], lE, t);
				}
				x = O::addmax(x, lE, ml);
				m = O::max3(d, x, y);
				Y[c] = y;
			}
			if (METHOD == M_SW) {
				if (c & 1)
					best = O::max3(best, mprev, m);
				else
					mprev = m;
			}
			ml = O::addw(m, dA, one); // fma pipe
			dg = t;
			Mo[c] = ml;
		}
		if (!last)
			bnd[r * 32] = make_uint2(ml, x);
	};

	int r = 0;
	uint32_t w = 0;
	uint2 b0 = make_uint2(0, 0), b1 = b0, b2 = b0, b3 = b0;
	w = rowp[0];
	if (!first) {
		b0 = bnd[0];
		b1 = bnd[32];
		b2 = bnd[64];
		b3 = bnd[96];
	}
	for (; r + 4 <= Lq; r += 4) {
		// prefetch the next 4-row group (buffers are over-allocated by one group)
		const uint32_t wn = rowp[((r >> 2) + 1) * 32];
		uint2 n0 = b0, n1 = b0, n2 = b0, n3 = b0;
		if (!first) {
			n0 = bnd[(r + 4) * 32];
			n1 = bnd[(r + 5) * 32];
			n2 = bnd[(r + 6) * 32];
			n3 = bnd[(r + 7) * 32];
		}
		step(w & 0xffu, b0, r);
		step((w >> 8) & 0xffu, b1, r + 1);
		step((w >> 16) & 0xffu, b2, r + 2);
		step(w >> 24, b3, r + 3);
		w = wn;
		b0 = n0;
		b1 = n1;
		b2 = n2;
		b3 = n3;
	}
	if (r < Lq) { // 1..3 tail rows
		step(w & 0xffu, b0, r);
		if (r + 1 < Lq) {
			step((w >> 8) & 0xffu, b1, r + 1);
			if (r + 2 < Lq)
				step((w >> 16) & 0xffu, b2, r + 2);
		}
	}

	// Global methods: the score is M[Lq][Ls] -> pick it out of the register strip when
	// the subject's last column lives in this strip (cap = Ls - 1 - c0, block-uniform).
	if (METHOD != M_SW) {
		if (cap0 >= 0 && cap0 < K) {
			uint32_t v = 0;
#pragma unroll
			for (int c = 0; c < K; c++)
				if (c == cap0)
					v = Mo[c];
			res = PACKED ? ((res & 0xffff0000u) | (v & 0x0000ffffu)) : v;
		}
		if (PACKED && cap1 >= 0 && cap1 < K) {
			uint32_t v = 0;
#pragma unroll
			for (int c = 0; c < K; c++)
				if (c == cap1)
					v = Mo[c];
			res = (res & 0x0000ffffu) | (v & 0xffff0000u);
		}
	}
}

// ---------------------------------------------------------------------------------------
// Panel-tile kernel: persistent blocks pull (query block x subject chunk) work items from
// an atomic counter, longest items first.
// ---------------------------------------------------------------------------------------
template <int METHOD, bool PACKED>
__global__ void __launch_bounds__(kThreads) sab_tile_kernel(const TileParams p)
{
	extern __shared__ uint32_t smem[];
	uint32_t *prof = smem; // [2][kWindow][kProfStride]
	int32_t *ssub = reinterpret_cast<int32_t *>(prof + 2 * kWindow * kProfStride); // [25][25]
	int *s_item = reinterpret_cast<int *>(ssub + kLetters * kLetters);

	constexpr int NS = PACKED ? 2 : 1;
	const int tid = threadIdx.x;
	const int lane = tid & 31;
	const int warp = tid >> 5;

	for (int i = tid; i < kLetters * kLetters; i += kThreads)
		ssub[i] = p.sub25[i];

	uint2 *const bnd = p.bnd + (int64_t)blockIdx.x * p.bnd_block_stride +
			   (int64_t)warp * p.bnd_rows * 32 + lane;
	const int a_end = p.a_slot0 + p.a_n;
	const int nitems = p.nqb * p.nsc;

	// Stage the pair profile for subjects (s, s+1), window w, into buffer `buf`.
	auto build = [&](const int s, const int s_end, const int w, const int buf) {
		if (tid < kWindow) {
			const int c = w * kWindow + tid;
			const int l0 = p.seqs.slot_len[s];
			int a0 = 24, a1 = 24;
			if (c < l0)
				a0 = p.seqs.col_bytes[p.seqs.col_off[s] + c];
			if (PACKED && s + 1 < s_end) {
				const int l1 = p.seqs.slot_len[s + 1];
				if (c < l1)
					a1 = p.seqs.col_bytes[p.seqs.col_off[s + 1] + c];
			}
			uint32_t *dst = prof + (buf * kWindow + tid) * kProfStride;
			const int32_t *r0 = ssub + a0 * kLetters;
			const int32_t *r1 = ssub + a1 * kLetters;
#pragma unroll 5
			for (int q = 0; q < kLetters; q++) {
				dst[q] = VOps<PACKED>::delta2(r0[q], r1[q]);
			}
		}
	};

	for (;;) {
		__syncthreads(); // previous item fully done (smem + s_item reusable)
		if (tid == 0)
			*s_item = (int)atomicAdd(p.counter, 1u);
		__syncthreads();
		const int item = *s_item;
		if (item >= nitems)
			break;
		// longest first: slots are sorted by ascending length inside a panel
		const int qb = p.nqb - 1 - item / p.nsc;
		const int sc = p.nsc - 1 - item % p.nsc;
		const int t0 = p.a_slot0 + qb * kThreads;
		const int t = t0 + tid; // my query slot
		const bool valid = t < a_end;
		int s_begin = sc * p.chunk;
		int s_end = s_begin + p.chunk;
		s_end = min(s_end, a_end);
		s_end = min(s_end, min(t0 + kThreads, a_end) - 1); // only subjects below the top query
		if (s_begin >= s_end)
			continue;

		const int Lq = valid ? p.seqs.slot_len[t] : 0;
		const int qorig = valid ? p.seqs.slot_orig[t] : 0;
		const uint32_t *rowp =
			p.seqs.rowpack + (valid ? p.seqs.grp_off[t >> 5] : 0) * 32 + lane;

		auto pair_len = [&](const int s) {
			int l = p.seqs.slot_len[s];
			if (PACKED && s + 1 < s_end)
				l = max(l, p.seqs.slot_len[s + 1]);
			return l;
		};

		int s = s_begin, w = 0, buf = 0;
		int Lp = pair_len(s);
		build(s, s_end, w, buf);
		__syncthreads();
		uint32_t best = 0, res = 0;
		const uint32_t one = p.one;
		while (s < s_end) {
			// next unit
			int ns = s, nw = w + 1, nLp = Lp;
			if (nw * kWindow >= Lp) {
				ns = s + NS;
				nw = 0;
				if (ns < s_end)
					nLp = pair_len(ns);
			}
			if (ns < s_end)
				build(ns, s_end, nw, buf ^ 1);

			const bool needed = valid && s < t;
			if (needed) {
				const int l0 = p.seqs.slot_len[s];
				const int l1 = (PACKED && s + 1 < s_end) ? p.seqs.slot_len[s + 1] : 0;
				if (w == 0) {
					best = VOps<PACKED>::bcast(0);
					res = 0;
				}
				const int cend = min((w + 1) * kWindow, Lp);
				const uint32_t *pw = prof + buf * kWindow * kProfStride;
				for (int c0 = w * kWindow; c0 < cend; c0 += kStrip) {
					run_strip<METHOD, PACKED>(pw + (c0 - w * kWindow) * kProfStride, rowp, bnd,
								  Lq, c0, c0 == 0, c0 + kStrip >= Lp, p.gapA,
								  p.gapE, one, best, res, l0 - 1 - c0,
								  l1 - 1 - c0);
				}
				if (cend == Lp) { // pair finished: emit scores
					using O = VOps<PACKED>;
					int sc0, sc1;
					if (METHOD == M_SW) {
						sc0 = O::lo(best);
						sc1 = O::hi(best);
					} else {
						sc0 = O::lo(res) - p.gapA;
						sc1 = O::hi(res) - p.gapA;
					}
					// a wrapped LOW lane carries into the high lane of the 32-bit adds:
					// its partner is re-run with it
					const bool lo_ovf = METHOD == M_SW && PACKED && sc0 > p.ovf_threshold;
#pragma unroll
					for (int h = 0; h < NS; h++) {
						const int ss = s + h;
						if (ss >= s_end || ss >= t)
							break;
						const int score = h ? sc1 : sc0;
						const int sorig = p.seqs.slot_orig[ss];
						const int hi = max(qorig, sorig), lo = min(qorig, sorig);
						p.blk[(int64_t)hi * (hi - 1) / 2 - p.tri0 + lo] = score;
						if (p.blkT)
							p.blkT[(int64_t)lo * p.a_n + (hi - p.a_orig0)] = score;
						if (METHOD == M_SW && PACKED && (score > p.ovf_threshold || lo_ovf)) {
							const unsigned k = atomicAdd(p.rerun_count, 1u);
							if (k < p.rerun_cap)
								p.rerun_list[k] =
									((unsigned long long)(unsigned)t << 32) | (unsigned)ss;
						}
					}
				}
			}
			__syncthreads(); // everyone done with `buf`; next profile visible
			s = ns;
			w = nw;
			Lp = nLp;
			buf ^= 1;
		}
	}
}

// ---------------------------------------------------------------------------------------
// int32 re-run of individual pairs whose int16 SW score may have wrapped.  One thread per
// queued pair, two rolling rows (M and Y) in a per-thread global scratch.  Rare path
// (needs a score > 32767 - max(S)); correctness over speed.
// ---------------------------------------------------------------------------------------
struct RerunParams {
	DeviceSeqs seqs;
	const unsigned long long *list;
	const unsigned *count;
	unsigned cap;
	int32_t *scratch; // [threads][2][scratch_cols]
	int scratch_cols;
	const int32_t *sub25; // as TileParams (offset by -gapA)
	int gapA, gapE;
	int a_orig0, a_n;
	int32_t *blk;
	int64_t tri0;
	int32_t *blkT;
};

__global__ void sab_rerun_sw_kernel(const RerunParams p)
{
	const unsigned n = min(*p.count, p.cap);
	const unsigned nthreads = gridDim.x * blockDim.x;
	const unsigned me = blockIdx.x * blockDim.x + threadIdx.x;
	int32_t *m = p.scratch + (int64_t)me * 2 * p.scratch_cols;
	int32_t *y = m + p.scratch_cols;
	const int o = p.gapA, e = p.gapE;
	const int NEG = INT32_MIN / 2;
	for (unsigned k = me; k < n; k += nthreads) {
		const unsigned long long ent = p.list[k];
		const int t = (int)(ent >> 32), s = (int)(ent & 0xffffffffu);
		const int Lq = p.seqs.slot_len[t], Ls = p.seqs.slot_len[s];
		const uint8_t *qs = p.seqs.col_bytes + p.seqs.col_off[t];
		const uint8_t *ss = p.seqs.col_bytes + p.seqs.col_off[s];
		for (int c = 0; c <= Ls; c++) {
			m[c] = 0;
			y[c] = NEG;
		}
		int best = 0;
		for (int r = 0; r < Lq; r++) {
			const int32_t *srow = p.sub25 + qs[r] * kLetters;
			int diag = m[0], left = 0, x = NEG;
			for (int c = 1; c <= Ls; c++) {
				const int sub = srow[ss[c - 1]] + o; // sub25 carries -gapA
				const int d = diag + sub;
				x = max(left + o, x + e);
				const int yy = max(m[c] + o, y[c] + e);
				int v = max(d, 0);
				v = max(v, x);
				v = max(v, yy);
				diag = m[c];
				m[c] = v;
				y[c] = yy;
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

} // namespace sab
