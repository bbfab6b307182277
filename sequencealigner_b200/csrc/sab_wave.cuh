// sab_wave.cuh -- systolic strip pipeline: the sm_100a kernel of the all-vs-all DP fill.
//
// Decomposition
//   * lane = QUERY sequence (DP rows, streamed 8 letters per step from the lane-interleaved
//     `rowpack`), block = 32 queries x a chunk of SUBJECT PAIRS (DP columns): the two
//     subjects of a pair sit in the low/high int16 halves of every register, so each DPX
//     instruction updates two cells.
//   * The 4 warps of a block work on the SAME 32 queries and on 4 CONSECUTIVE K-column
//     strips (K = 32 or 16) of the stream of strips of the item's subject pairs:
//            step tau:   warp j computes row group (tau - j) mod G of its strip
//     so warp j+1 is always one row group (8 rows) behind warp j, and the strip boundary
//     column (V, X per row) travels warp j -> warp j+1 through a 12 KB shared-memory
//     exchange buffer (the role the north star gives to __shfl_sync; lanes cannot be used
//     for it because the conflict-free profile read needs all lanes of a warp on the same
//     DP column).  Only the wrap-around warp 3 -> warp 0 (every 4K columns) goes through a
//     per-block global buffer small enough to stay in L2.  The stream never drains inside
//     an item: a warp that finishes a strip starts the strip four positions further on.
//   * Each strip's substitution scores are expanded once into a shared-memory PAIR PROFILE
//     prof[column][query letter] = (S[a0_col][q], S[a1_col][q]) - shift.  All 32 lanes read
//     the same column with their own letter: 25 consecutive words, never a bank conflict.
//
// Cell arithmetic (per packed cell pair; alu = DPX on the 16-lane alu pipe, fma = IMAD)
//   The stored matrix value V is a gap-skewed image of the reference's M/H so that the
//   recurrences need as few adds as possible (g, o, e <= 0 are GAP_PEN/GAP_OPN/GAP_EXT):
//     NW  V = H - g(r+c)      V = max3(Vdiag + (S - 2g), Vleft, Vup)               1 alu 1 fma
//     GA  V = M - e(r+c)      X = max(Vleft + (o-e), Xleft)   Y = max(Vup + (o-e), Yup)
//         (X, Y skewed too)   V = max3(Vdiag + (S - 2e), X, Y)                      3 alu 1 fma
//     SW  V = M + o           D = max(Vdiag + (S - o), 0)  Y = max(Yup + e, Vup)  X = max(Xleft + e, Vleft)
//         (floor on D only)   M = max3(D, X, Y)   V = M + o   best = max3(best, M, M')   4.5 alu 1 fma
//   NW/GA borders become constants (0 resp. o-e) and the score is V + g|e * (Lq + Ls).
//   SW keeps the unskewed form because its zero floor would turn into a per-cell value.
//   An 8-row group is issued in ANTI-DIAGONAL order (row k works on column t-k at time t):
//   the 2-3 instruction dependency chain of a cell is interleaved with seven independent
//   ones, which is what lets 4-5 resident warps per scheduler keep the alu pipe busy
//   (profiles/r1_ncu_wave_v1.md shows the serial order stalling at 65 % alu; now GA 78 %,
//   SW 88 %: r1_ncu_wave_v5_*.md; what this code shape can reach: r1_cell_shape_ceiling.md).
//
// One __syncthreads per step (8 rows x K columns x 2 cells per thread).
#pragma once

#include "sab_kernels.cuh"

#include <type_traits>

namespace sab {

#ifndef SAB_KGROUP
#define SAB_KGROUP 8
#endif
constexpr int kGroup = SAB_KGROUP; // rows per step (multiple of 4: rowpack holds 4 letters per word)
constexpr int kGroupWords = kGroup / 4;
constexpr int kBestSlots = 16; // subject pairs whose SW maxima can be in flight
// Steps per round: round r+1 is staged at step 3 of round r, the earliest step at which warp 3 has
// left round r-1 (whose profile buffer it overwrites), and must be done before warp 0 enters round
// r+1 at step G; the wrap-around boundary written by warp 3 at step g+3 is read by warp 0 G-3 steps
// later.  Both need G >= 4.
constexpr int kMinSteps = 4;

template <int K> struct WaveSmem {
	uint32_t prof[2][kWaveWarps][K][kProfStride]; // strip profiles, parity = round & 1
	uint2 xch[2][kWaveWarps - 1][kGroup][32]; // boundary rows warp j -> j+1, parity = step & 1
	uint32_t sbest[kBestSlots][32]; // SW running maxima per subject pair
	int32_t ssub[kLetters * kLetters];
	int desc[2][kWaveWarps][8]; // strip descriptors, parity = round & 1
	int item;
	uint2 cbnd[kGroup][32]; // kHoistExchange: the boundary rows a pair's FIRST strip starts from, (border V, -inf)
};

// Hoisted boundary exchange (NW and GA).  Left to itself the compiler re-derives the shared-window address of
// the exchange buffer in every step of every warp -- S2R SR_CgaCtaId, LEA and a 6-instruction chain in front of
// the boundary loads, again in front of the stores -- executes the 16 register moves of the "first strip"
// constants whether the strip is a first one or not, and rebuilds the 64-bit address of the row letters from
// g.  Here the three 32-bit shared addresses are formed once per kernel and made opaque (an empty asm), the
// accesses are ld/st.shared with immediate row offsets, a pair's first strip reads its constant boundary from
// shared memory (cbnd) like any other, the parity of the exchange buffer is an XOR on a uniform register and
// the row pointer runs along.  ~45 of the ~165 bookkeeping instructions of a step and the S2R latency in front
// of its first cell go away: C2 8.99 -> 9.16 TCUPS, C4 6.78 -> 7.17, same box (profiles/r2c_ab_hoist.txt).
// SW got 0.9 % slower with it (three more live registers at 123 of 128): off there.
#ifndef SAB_HOIST_METHODS
#define SAB_HOIST_METHODS(method) ((method) != sab::M_SW)
#endif
template <int OFF> __device__ __forceinline__ uint2 lds64(const uint32_t a)
{
	uint2 v;
	asm volatile("ld.shared.v2.u32 {%0, %1}, [%2+%3];" : "=r"(v.x), "=r"(v.y) : "r"(a), "n"(OFF) : "memory");
	return v;
}
template <int OFF> __device__ __forceinline__ void sts64(const uint32_t a, const uint32_t x, const uint32_t y)
{
	asm volatile("st.shared.v2.u32 [%0+%1], {%2, %3};" ::"r"(a), "n"(OFF), "r"(x), "r"(y) : "memory");
}
constexpr int kXchRowBytes = 32 * sizeof(uint2); // one row of the exchange buffer: a uint2 per lane
// rows k = 0 .. NR-1 of my lane, starting at shared address a
template <int NR, int k = 0> __device__ __forceinline__ void lds_rows(const uint32_t a, uint2 (&b)[NR])
{
	if constexpr (k < NR) {
		b[k] = lds64<k * kXchRowBytes>(a);
		lds_rows<NR, k + 1>(a, b);
	}
}
template <int NR, int k = 0>
__device__ __forceinline__ void sts_rows(const uint32_t a, const uint32_t (&v)[NR], const uint32_t (&x)[NR])
{
	if constexpr (k < NR) {
		sts64<k * kXchRowBytes>(a, v[k], x[k]);
		sts_rows<NR, k + 1>(a, v, x);
	}
}

// Per-launch constants of the cell update, in the lane format of VOps<PACKED>.
template <int METHOD, bool PACKED> struct CellK {
	uint32_t one; // IMAD multiplier (uniform register)
	uint32_t lG; // GA: lane(o - e); SW: lane(e)
	uint32_t dE; // SW: 32-bit-add form of "+ e"
	uint32_t dA; // SW: 32-bit-add form of "+ o"
	uint32_t ZERO; // SW floor
	int gE; // SW int32: raw extend
};

// One DP cell.  V = stored value of this column (in: row above, out: this row),
// Yc = vertical-gap state of the column, ml/x = left neighbour's V / horizontal-gap state,
// dg = diagonal neighbour's V, s = profile entry.  Returns M of the cell (SW maximum).
template <int METHOD, bool PACKED>
__device__ __forceinline__ uint32_t cell(const CellK<METHOD, PACKED> &k, uint32_t &V, uint32_t &Yc,
					 uint32_t &ml, uint32_t &x, uint32_t &dg, const uint32_t s)
{
	using O = VOps<PACKED>;
	const uint32_t tt = V;
	const uint32_t d = (METHOD == M_SW && PACKED) ? 0u : O::addw(dg, s, k.one); // fma pipe
	uint32_t m;
	if (METHOD == M_NW) {
		m = O::max3(d, ml, tt);
		ml = m;
	} else if (METHOD == M_GA) {
		const uint32_t y = O::addmax(tt, k.lG, Yc);
		x = O::addmax(ml, k.lG, x);
		m = O::max3(d, x, y);
		Yc = y;
		ml = m;
	} else {
		uint32_t y;
		if (PACKED) {
			// The zero floor sits on the diagonal term (M = max(0, Mdiag + S, X, Y); a negative
			// X or Y can never win against it nor against a later M + o), so both gap adds and
			// the score add ride inside DPX add-max ops: 4.5 alu + 1 fma per cell pair.  Measured
			// (tools/cell_peak.cu) 4 % faster than 3.5 alu + 3 fma: an IMAD next to a saturated
			// DPX stream costs ~0.8 clk, not 0.  `s` is in per-lane encoding here (O::lane2).
			const uint32_t d2 = O::addmax(dg, s, k.ZERO);
			y = O::addmax(Yc, k.lG, tt);
			x = O::addmax(x, k.lG, ml);
			m = O::max3(d2, x, y);
			Yc = y;
		} else {
			y = (uint32_t)__viaddmax_s32_relu((int)Yc, k.gE, (int)tt);
			Yc = y;
			x = O::addmax(x, k.lG, ml);
			m = O::max3(d, x, y);
		}
		ml = O::addw(m, k.dA, k.one); // fma pipe
	}
	dg = tt;
	V = ml;
	return m;
}

#ifdef SAB_PICK_SWITCH
// experiment: V[idx] for a block-uniform idx as a jump table instead of K compare/select pairs
template <int K> __device__ __forceinline__ uint32_t pick_column(const uint32_t (&V)[K], const int idx)
{
	switch (idx) {
#define SAB_PICK_CASE(i) \
	case i:          \
		return V[i < K ? i : 0];
		SAB_PICK_CASE(0) SAB_PICK_CASE(1) SAB_PICK_CASE(2) SAB_PICK_CASE(3) SAB_PICK_CASE(4) SAB_PICK_CASE(5)
		SAB_PICK_CASE(6) SAB_PICK_CASE(7) SAB_PICK_CASE(8) SAB_PICK_CASE(9) SAB_PICK_CASE(10) SAB_PICK_CASE(11)
		SAB_PICK_CASE(12) SAB_PICK_CASE(13) SAB_PICK_CASE(14) SAB_PICK_CASE(15) SAB_PICK_CASE(16) SAB_PICK_CASE(17)
		SAB_PICK_CASE(18) SAB_PICK_CASE(19) SAB_PICK_CASE(20) SAB_PICK_CASE(21) SAB_PICK_CASE(22) SAB_PICK_CASE(23)
		SAB_PICK_CASE(24) SAB_PICK_CASE(25) SAB_PICK_CASE(26) SAB_PICK_CASE(27) SAB_PICK_CASE(28) SAB_PICK_CASE(29)
		SAB_PICK_CASE(30) SAB_PICK_CASE(31)
#undef SAB_PICK_CASE
	}
	return 0;
}
#endif

// K = DP columns a warp holds in registers (strip width): 16..32 in steps of 2, chosen per row
// strip by the host (sab_capi.cu: choose_widths); narrow strips fit 5 resident blocks per SM.
template <int METHOD, bool PACKED, int K, int MINB>
__global__ void __launch_bounds__(kWaveThreads, MINB) sab_wave_kernel(const TileParams p)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	WaveSmem<K> &sm = *reinterpret_cast<WaveSmem<K> *>(smem_raw);
	using O = VOps<PACKED>;
	constexpr int NS = PACKED ? 2 : 1;
#ifdef SAB_PREFETCH_ROWS
	constexpr bool kPrefetchRows = SAB_PREFETCH_ROWS;
#else
	constexpr bool kPrefetchRows = false;
#endif
	const int tid = threadIdx.x;
	const int lane = tid & 31;
	const int warp = tid >> 5;

	for (int i = tid; i < kLetters * kLetters; i += kWaveThreads)
		sm.ssub[i] = p.sub25[i];

	// wrap-around boundary (warp 3 -> warp 0 of the next round): [row][lane]
	uint2 *const gfr = p.bnd + (int64_t)blockIdx.x * p.bnd_block_stride + lane;
	const int a_end = p.a_slot0 + p.a_n;
	const int nqg = (p.a_n + 31) >> 5;
	const int nitems = nqg * p.nsc;
	const int gA = p.gapA, gE = p.gapE;

	CellK<METHOD, PACKED> ck;
	ck.one = p.one;
#ifdef SAB_LG_IMM
	// Product build: the per-lane gap constant of the two add-max ops of every cell is an IMMEDIATE
	// operand (one register-file read less per DPX op: C2 +3.2 %, C3 +5.4 %, profiles/r2_ab_call1.txt).
	// The image carries the placeholder SAB_LG_IMM; sab_module.cpp writes the job's value into the
	// instruction words before the image is loaded.
	ck.lG = SAB_LG_IMM;
#else
	ck.lG = O::lane(METHOD == M_GA ? gA - gE : gE);
#endif
	ck.dE = O::delta(gE);
	ck.dA = O::delta(gA);
	ck.ZERO = O::bcast(0);
	ck.gE = gE;
	// -inf stand-in.  NW/GA never add to it; SW adds e once, so NEG + e == biased 0 there.
	const uint32_t NEG = PACKED ? (METHOD == M_SW ? O::lane(-gE) : 0u) : O::neg_inf(gE);
	// Borders in stored coordinates (header comment): every border cell holds the same value.
	//   NW  H[0][c] = cg, H[r][0] = rg                          -> V = 0        (nw.c:13-17)
	//   GA  M[0][c] = o+(c-1)e, M[r][0] = o+(r-1)e, M[0][0] = 0 -> V = o-e, 0   (ga.c:23-38)
	//   SW  M = 0 on both borders                               -> V = o        (sw.c:18-30)
	const int vb = METHOD == M_NW ? 0 : METHOD == M_GA ? gA - gE : gA;
	const int v00 = METHOD == M_SW ? gA : 0;
	const uint32_t VB = O::bcast(vb);
	const int gS = METHOD == M_NW ? gA : gE; // score = V + gS * (Lq + Ls) for NW/GA
	constexpr bool kHoistExchange = SAB_HOIST_METHODS(METHOD);
	constexpr uint32_t kXchWarp = kGroup * 32 * sizeof(uint2); // one [warp] slab of xch
	constexpr uint32_t kXchPar = (kWaveWarps - 1) * kXchWarp; // one parity of xch
	uint32_t xin0 = 0, xout0 = 0, s_cbnd = 0; // shared-window addresses of my lane's first rows
	if constexpr (kHoistExchange) {
		const uint32_t s_xch =
			(uint32_t)__cvta_generic_to_shared(&sm.xch[0][0][0][0]) + lane * (uint32_t)sizeof(uint2);
		xin0 = s_xch + (warp ? warp - 1 : 0) * kXchWarp;
		xout0 = s_xch + (warp < kWaveWarps - 1 ? warp : 0) * kXchWarp;
		s_cbnd = (uint32_t)__cvta_generic_to_shared(&sm.cbnd[0][0]) + lane * (uint32_t)sizeof(uint2);
		// opaque to the compiler, or it re-derives them from SR_CgaCtaId at every use (cheaper by its cost model)
		asm volatile("" : "+r"(xin0), "+r"(xout0), "+r"(s_cbnd));
		for (int i = tid; i < kGroup * 32; i += kWaveThreads)
			(&sm.cbnd[0][0])[i] = make_uint2(VB, NEG); // visible after the item loop's first barrier
	}

	for (;;) {
		__syncthreads(); // previous item fully done
		if (tid == 0)
			sm.item = (int)atomicAdd(p.counter, 1u);
		__syncthreads();
		const int item = sm.item;
		if (item >= nitems)
			break;
		// longest first: slots are sorted by ascending length inside a panel
		const int qg = nqg - 1 - item / p.nsc;
		const int ck_i = p.nsc - 1 - item % p.nsc;
		const int sc = p.chunk_ids ? p.chunk_ids[ck_i] : ck_i;
		const int t0 = p.a_slot0 + qg * 32;
		const int t = t0 + lane; // my query slot
		const bool valid = t < a_end;
		const int s_begin = sc * p.chunk;
		int s_end = min(s_begin + p.chunk, a_end);
		s_end = min(s_end, min(t0 + 32, a_end) - 1); // only subjects below the top query
		if (s_begin >= s_end)
			continue;

		const int Lq = valid ? p.seqs.slot_len[t] : 0;
		const int qorig = valid ? p.seqs.slot_orig[t] : 0;
		const uint32_t *rowp = p.seqs.rowpack + p.seqs.grp_off[t0 >> 5] * 32 + lane;
		const int Lmax = __reduce_max_sync(0xffffffffu, Lq);
		const int G = max(kMinSteps, (Lmax + kGroup - 1) / kGroup); // steps per round

		auto pair_len = [&](const int s) {
			int l = p.seqs.slot_len[s];
			if (PACKED && s + 1 < s_end)
				l = max(l, p.seqs.slot_len[s + 1]);
			return l;
		};
		// rounds of the item: 4 strips of the stream per round
		int total_strips = 0;
		for (int s = s_begin; s < s_end; s += NS)
			total_strips += (pair_len(s) + K - 1) / K;
		const int R = (total_strips + kWaveWarps - 1) / kWaveWarps;

		// strip stream iterator (block uniform)
		int it_s = s_begin, it_c = 0, it_Lp = pair_len(s_begin), it_pair = 0;

		// Stages round `rnd`: strip descriptors + the 32-column profile of each of its strips.
		auto build = [&](const int rnd) {
			const int pb = rnd & 1;
			int my_s = -1, my_c0 = 0, my_l0 = 0, my_l1 = 0, my_Lp = 0, my_ps = 0;
#pragma unroll
			for (int jj = 0; jj < kWaveWarps; jj++) {
				if (it_s < s_end) {
					if (jj == warp) {
						my_s = it_s;
						my_c0 = it_c;
						my_Lp = it_Lp;
						my_ps = it_pair & (kBestSlots - 1);
						my_l0 = p.seqs.slot_len[it_s];
						my_l1 = (PACKED && it_s + 1 < s_end) ? p.seqs.slot_len[it_s + 1] : 0;
					}
					it_c += K;
					if (it_c >= it_Lp) {
						it_s += NS;
						it_c = 0;
						it_pair++;
						if (it_s < s_end)
							it_Lp = pair_len(it_s);
					}
				}
			}
			if (lane == 0) {
				int *d = sm.desc[pb][warp];
				d[0] = my_s;
				d[1] = my_c0;
				d[2] = my_l0;
				d[3] = my_l1;
				d[4] = my_Lp;
				d[5] = my_ps;
			}
			int a0 = 24, a1 = 24;
			if (my_s >= 0) {
				const int c = my_c0 + lane;
				if (c < my_l0)
					a0 = subject_letter(p.seqs, my_s, c);
				if (PACKED && c < my_l1)
					a1 = subject_letter(p.seqs, my_s + 1, c);
			}
			if (K == 32 || lane < K) { // lane = column of the strip
				uint32_t *dst = sm.prof[pb][warp][lane < K ? lane : 0];
				const int32_t *r0 = sm.ssub + a0 * kLetters;
				const int32_t *r1 = sm.ssub + a1 * kLetters;
#pragma unroll 5
				for (int q = 0; q < kLetters; q++)
					dst[q] = METHOD == M_SW ? O::lane2(r0[q], r1[q]) : O::delta2(r0[q], r1[q]);
			}
		};

		build(0);
		__syncthreads();

		// per-warp pipeline position and strip state
		int g = 0, rho = 0, u = -warp;
		int tau_mod = 0, tau_round = 0;
		int d_s = -1, d_c0 = 0, d_l0 = 0, d_l1 = 0, d_ps = 0;
		bool first = false, last = false;
		uint32_t V[K], Y[K];
		// experiment (-DSAB_PREFETCH_ROWS=1): the next row group's letters are loaded one step ahead.
		// SW + 2.8 % on C3 before the gap constant became an immediate, - 1.4 % after; GA +- 0.5 %: off.
		uint32_t wn[kGroupWords];
		if (kPrefetchRows) {
#pragma unroll
			for (int i = 0; i < kGroupWords; i++)
				wn[i] = rowp[i * 32];
		}
		uint32_t diag = 0, bestA = ck.ZERO, bestB = ck.ZERO;
		const uint32_t *prof_s = sm.prof[0][warp][0];
		const int T = R * G + (kWaveWarps - 1);
		uint32_t xpar = 0; // kHoistExchange: (tau & 1) * kXchPar
		const uint32_t *rowg = rowp; // kHoistExchange: rowp + g * kGroupWords * 32

		for (int tau = 0; tau < T; tau++) {
			if (tau_mod == 3 && tau_round + 1 < R)
				build(tau_round + 1);

			if (u >= 0 && rho < R) {
				if (g == 0) { // start of a strip: DP row 0 in stored coordinates
					const int *d = sm.desc[rho & 1][warp];
					d_s = d[0];
					d_c0 = d[1];
					d_l0 = d[2];
					d_l1 = d[3];
					const int d_Lp = d[4];
					d_ps = d[5];
					first = d_c0 == 0;
					last = d_c0 + K >= d_Lp;
					prof_s = sm.prof[rho & 1][warp][0];
					const uint32_t y0 = (METHOD == M_SW && PACKED) ? ck.ZERO : NEG;
#pragma unroll
					for (int c = 0; c < K; c++) {
						V[c] = VB;
						Y[c] = y0;
					}
					diag = first ? O::bcast(v00) : VB;
					bestA = bestB = ck.ZERO;
					if (METHOD == M_SW && first && d_s >= 0)
						sm.sbest[d_ps][lane] = ck.ZERO;
				}
				const int r0 = g * kGroup;
				uint32_t w[kGroupWords];
				if (kPrefetchRows) {
					const int gn = g + 1 == G ? 0 : g + 1;
#pragma unroll
					for (int i = 0; i < kGroupWords; i++) {
						w[i] = wn[i];
						if (gn * kGroup < Lmax)
							wn[i] = rowp[(gn * kGroupWords + i) * 32];
					}
				}
				if (d_s >= 0 && r0 < Lq) {
					if (!kPrefetchRows) {
#pragma unroll
						for (int i = 0; i < kGroupWords; i++)
							w[i] = kHoistExchange ? rowg[i * 32] : rowp[(g * kGroupWords + i) * 32];
					}
					const uint32_t xout_a = xout0 + xpar;
					const uint2 *xin = &sm.xch[(tau - 1) & 1][warp ? warp - 1 : 0][0][lane];
					uint2 *xout = &sm.xch[tau & 1][warp < kWaveWarps - 1 ? warp : 0][0][lane];
					const bool from_g = warp == 0, to_g = warp == kWaveWarps - 1;
					// NR rows [k0, k0 + NR) of the group as straight-line code, issued
					// anti-diagonally: at time tt row k works on column tt - k.
					auto run_rows = [&](auto nr_c, const int k0) {
						constexpr int NR = decltype(nr_c)::value;
						uint2 b[NR];
						if constexpr (kHoistExchange) {
							if (from_g && !first) {
#pragma unroll
								for (int k = 0; k < NR; k++)
									b[k] = gfr[(r0 + k0 + k) * 32];
							} else {
								// a pair's first strip reads the constant rows, every other one
								// what the warp to its left wrote in the previous step
								lds_rows<NR>((first ? s_cbnd : xin0 + (xpar ^ kXchPar)) + k0 * kXchRowBytes, b);
							}
						} else if (first) {
#pragma unroll
							for (int k = 0; k < NR; k++)
								b[k] = make_uint2(VB, NEG);
						} else if (from_g) {
#pragma unroll
							for (int k = 0; k < NR; k++)
								b[k] = gfr[(r0 + k0 + k) * 32];
						} else {
#pragma unroll
							for (int k = 0; k < NR; k++)
								b[k] = xin[(k0 + k) * 32];
						}
						uint32_t ml[NR], x[NR], dg[NR], mprev[NR];
						const uint32_t *pq[NR];
#pragma unroll
						for (int k = 0; k < NR; k++) {
							ml[k] = b[k].x;
							x[k] = b[k].y;
							dg[k] = k ? b[k - 1].x : diag;
							mprev[k] = 0;
							const int kk = k0 + k;
							const uint32_t wk = (kGroupWords > 1 && (kk & 4)) ? w[kGroupWords - 1] : w[0];
							pq[k] = reinterpret_cast<const uint32_t *>(
								reinterpret_cast<const char *>(prof_s) + ((wk >> (8 * (kk & 3))) & 0xffu));
						}
						diag = b[NR - 1].x;
#pragma unroll
						for (int tt = 0; tt < K + NR - 1; tt++) {
#pragma unroll
							for (int k = 0; k < NR; k++) {
								const int c = tt - k;
								if (c < 0 || c >= K)
									continue;
								const uint32_t m = cell<METHOD, PACKED>(
									ck, V[c], Y[c], ml[k], x[k], dg[k], pq[k][c * kProfStride]);
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
						if (!last) {
							if (to_g) {
#pragma unroll
								for (int k = 0; k < NR; k++)
									gfr[(r0 + k0 + k) * 32] = make_uint2(ml[k], x[k]);
							} else {
								if constexpr (kHoistExchange) {
									sts_rows<NR>(xout_a + k0 * kXchRowBytes, ml, x);
								} else {
#pragma unroll
									for (int k = 0; k < NR; k++)
										xout[(k0 + k) * 32] = make_uint2(ml[k], x[k]);
								}
							}
						}
					};
					// SW runs its last group in full: the rows past Lq hold the pad letter
					// (score 0 against everything), which can never raise the running maximum.
					// NW/GA need V at row Lq exactly: their 1..7 tail rows go as 4 + 2 + 1.
					if (METHOD == M_SW || r0 + kGroup <= Lq) {
						run_rows(std::integral_constant<int, kGroup>{}, 0);
					} else {
						const int nv = Lq - r0;
						if (nv & 4)
							run_rows(std::integral_constant<int, 4>{}, 0);
						if (nv & 2)
							run_rows(std::integral_constant<int, 2>{}, nv & 4);
						if (nv & 1)
							run_rows(std::integral_constant<int, 1>{}, nv & 6);
					}
					// ---- my last row of this strip: results ------------------------
					if (Lq <= r0 + kGroup) {
						int sc0 = 0, sc1 = 0;
						bool have0 = false, have1 = false;
						if (METHOD == M_SW) {
							uint32_t acc = sm.sbest[d_ps][lane];
							acc = O::max3(acc, bestA, bestB);
							if (last) {
								sc0 = O::lo(acc);
								sc1 = O::hi(acc);
								have0 = true;
								have1 = PACKED;
							} else {
								sm.sbest[d_ps][lane] = acc;
							}
						} else {
							// M[Lq][Ls] = V + gS * (Lq + Ls), picked out of the register strip
							// that holds the subject's last column (block-uniform index)
							const int cap0 = d_l0 - 1 - d_c0, cap1 = d_l1 - 1 - d_c0;
							if (cap0 >= 0 && cap0 < K) {
								uint32_t v = 0;
#ifdef SAB_PICK_SWITCH
								v = pick_column<K>(V, cap0);
#else
#pragma unroll
								for (int c = 0; c < K; c++)
									if (c == cap0)
										v = V[c];
#endif
								sc0 = O::lo(v) + gS * (Lq + d_l0);
								have0 = true;
							}
							if (PACKED && cap1 >= 0 && cap1 < K) {
								uint32_t v = 0;
#ifdef SAB_PICK_SWITCH
								v = pick_column<K>(V, cap1);
#else
#pragma unroll
								for (int c = 0; c < K; c++)
									if (c == cap1)
										v = V[c];
#endif
								sc1 = O::hi(v) + gS * (Lq + d_l1);
								have1 = true;
							}
						}
						// a wrapped LOW lane carries into the high lane of the 32-bit adds:
						// its partner is re-run with it
						const bool lo_ovf = METHOD == M_SW && PACKED && sc0 > p.ovf_threshold;
#pragma unroll
						for (int h = 0; h < NS; h++) {
							const int ss = d_s + h;
							if (!(h ? have1 : have0) || ss >= s_end || ss >= t)
								continue;
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
				if (++g == G) {
					g = 0;
					rho++;
				}
				if (kHoistExchange)
					rowg = g ? rowg + kGroupWords * 32 : rowp;
			}
			if (kHoistExchange)
				xpar ^= kXchPar;
			u++;
			if (++tau_mod == G) {
				tau_mod = 0;
				tau_round++;
			}
			__syncthreads();
		}
	}
}

} // namespace sab
