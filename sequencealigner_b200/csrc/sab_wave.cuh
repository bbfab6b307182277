// sab_wave.cuh -- systolic strip pipeline: the main sm_100a kernel of the all-vs-all DP fill.
//
// Same cell arithmetic and data layouts as sab_kernels.cuh (biased int16x2 lanes, DPX ops
// on the alu pipe, carry-free 32-bit adds on the fma pipe, [column][letter] pair profile
// in shared memory), different decomposition.  The first kernel gave every thread a whole
// (query x subject pair) alignment and swept its 32-column register strips one after the
// other, parking the strip boundary column (M+open, X per row) in global memory.  ncu on
// C2 showed that boundary parking to be the limiter: 300 GB of DRAM traffic per launch and
// long-scoreboard stalls on the boundary loads (profiles/r1_ncu_panel_kernel.md).
//
// Here the 4 warps of a block work on the SAME 32 queries (lane = query) and on 4
// CONSECUTIVE strips of the stream of strips of the item's subject pairs:
//
//        step tau:   warp j computes row group (tau - j) mod G of its strip
//
// so warp j+1 is always one row group (4 rows) behind warp j, and the boundary column
// travels warp j -> warp j+1 through an 8 KB shared-memory exchange buffer (the role the
// north star gives to __shfl_sync; lanes cannot be used for it because the conflict-free
// profile read needs all lanes of a warp on the same DP column).  Only the wrap-around
// warp 3 -> warp 0 (every 128 columns) still goes through global memory, one round later,
// from a per-block buffer small enough to stay in L2.  The stream never drains inside an
// item: when a warp finishes a strip it starts the strip four positions further on, of the
// same or the next subject pair, so all four warps stay busy whatever the lengths are.
//
// One __syncthreads per step (4 rows x 32 columns x 2 cells per thread).
#pragma once

#include "sab_kernels.cuh"

namespace sab {

constexpr int kWaveWarps = 4;
constexpr int kWaveThreads = kWaveWarps * 32;
constexpr int kGroup = 4; // rows per step
constexpr int kBestSlots = 16; // subject pairs whose SW maxima can be in flight
constexpr int kMinSteps = 6; // steps per round (>= 4 for the double-buffered staging)

struct WaveSmem {
	uint32_t prof[2][kWaveWarps][kStrip][kProfStride]; // strip profiles, parity = round & 1
	uint2 xch[2][kWaveWarps - 1][kGroup][32]; // boundary rows warp j -> j+1, parity = step & 1
	uint32_t sbest[kBestSlots][32]; // SW running maxima per subject pair
	int32_t ssub[kLetters * kLetters];
	int desc[2][kWaveWarps][8]; // strip descriptors, parity = round & 1
	int item;
};
constexpr int kWaveSmemBytes = (int)sizeof(WaveSmem);

template <int METHOD, bool PACKED, int MINB>
__global__ void __launch_bounds__(kWaveThreads, MINB) sab_wave_kernel(const TileParams p)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	WaveSmem &sm = *reinterpret_cast<WaveSmem *>(smem_raw);
	using O = VOps<PACKED>;
	constexpr int NS = PACKED ? 2 : 1;
	constexpr int K = kStrip;
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
	const uint32_t one = p.one;
	const uint32_t dA = O::delta(gA);
	const uint32_t dE = O::delta(gE);
	const uint32_t lE = O::lane(gE);
	const uint32_t NEG = O::neg_inf(gE);
	const uint32_t ZERO = O::bcast(0);

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
		const int sc = p.nsc - 1 - item % p.nsc;
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
					a0 = p.seqs.col_bytes[p.seqs.col_off[my_s] + c];
				if (PACKED && c < my_l1)
					a1 = p.seqs.col_bytes[p.seqs.col_off[my_s + 1] + c];
			}
			uint32_t *dst = sm.prof[pb][warp][lane];
			const int32_t *r0 = sm.ssub + a0 * kLetters;
			const int32_t *r1 = sm.ssub + a1 * kLetters;
#pragma unroll 5
			for (int q = 0; q < kLetters; q++)
				dst[q] = O::delta2(r0[q], r1[q]);
		};

		build(0);
		__syncthreads();

		// per-warp pipeline position and strip state
		int g = 0, rho = 0, u = -warp;
		int tau_mod = 0, tau_round = 0;
		int d_s = -1, d_c0 = 0, d_l0 = 0, d_l1 = 0, d_ps = 0;
		bool first = false, last = false;
		uint32_t Mo[K], Y[K];
		uint32_t diag = 0, bl = 0, dstep = 0, best = ZERO;
		const uint32_t *prof_s = sm.prof[0][warp][0];
		const int T = R * G + (kWaveWarps - 1);

		for (int tau = 0; tau < T; tau++) {
			if (tau_mod == 3 && tau_round + 1 < R)
				build(tau_round + 1);

			if (u >= 0 && rho < R) {
				if (g == 0) { // start of a strip
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
					// Row 0 of the DP (borders), see run_strip() in sab_kernels.cuh
					int m0, st, d0, b0;
					if (METHOD == M_NW) {
						m0 = (d_c0 + 2) * gA;
						st = gA;
						d0 = (d_c0 + 1) * gA;
						b0 = 2 * gA;
					} else if (METHOD == M_GA) {
						m0 = 2 * gA + d_c0 * gE;
						st = gE;
						d0 = d_c0 ? 2 * gA + (d_c0 - 1) * gE : gA;
						b0 = 2 * gA;
					} else {
						m0 = gA;
						st = 0;
						d0 = gA;
						b0 = gA;
					}
					dstep = O::delta(st);
					Mo[0] = O::bcast(m0);
					Y[0] = (METHOD == M_SW && PACKED) ? ZERO : NEG;
#pragma unroll
					for (int c = 1; c < K; c++) {
						Mo[c] = Mo[c - 1] + dstep;
						Y[c] = Y[0];
					}
					diag = O::bcast(d0);
					bl = O::bcast(b0);
					best = ZERO;
					if (METHOD == M_SW && first && d_s >= 0)
						sm.sbest[d_ps][lane] = ZERO;
				}
				const int r0 = g * kGroup;
				if (d_s >= 0 && r0 < Lq) {
					// One DP row of the strip: boundary (ml, x) of the column to the left in,
					// boundary of the strip's last column out.
					auto row = [&](uint32_t ml, uint32_t x, const uint32_t qoff) -> uint2 {
						const uint32_t *pq = reinterpret_cast<const uint32_t *>(
							reinterpret_cast<const char *>(prof_s) + qoff);
						uint32_t dg = diag;
						diag = ml;
						uint32_t mprev = 0;
#pragma unroll
						for (int c = 0; c < K; c++) {
							const uint32_t s = pq[c * kProfStride];
							const uint32_t tt = Mo[c];
							const uint32_t d = O::addw(dg, s, one); // fma pipe
							uint32_t m;
							if (METHOD == M_NW) {
								m = O::max3(d, ml, tt);
							} else {
								uint32_t y;
								if (METHOD == M_SW) {
									if (PACKED) {
										const uint32_t ye = O::addw(Y[c], dE, one);
										y = O::max3(ye, tt, ZERO);
									} else {
										y = (uint32_t)__viaddmax_s32_relu((int)Y[c], gE, (int)tt);
									}
								} else {
									y = O::addmax(Y[c], lE, tt);
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
							dg = tt;
							Mo[c] = ml;
						}
						return make_uint2(ml, x);
					};
					const uint32_t w = rowp[g * 32];
					const uint2 *xin = &sm.xch[(tau - 1) & 1][warp ? warp - 1 : 0][0][lane];
					uint2 *xout = &sm.xch[tau & 1][warp < kWaveWarps - 1 ? warp : 0][0][lane];
					const bool from_g = warp == 0, to_g = warp == kWaveWarps - 1;
					if (r0 + kGroup <= Lq) {
						// ---- full group: straight-line code, no per-row control flow ----
						uint2 b[kGroup];
						if (first) {
#pragma unroll
							for (int k = 0; k < kGroup; k++) {
								b[k] = make_uint2(bl, NEG);
								bl += dstep;
							}
						} else if (from_g) {
#pragma unroll
							for (int k = 0; k < kGroup; k++)
								b[k] = gfr[(r0 + k) * 32];
						} else {
#pragma unroll
							for (int k = 0; k < kGroup; k++)
								b[k] = xin[k * 32];
						}
#pragma unroll
						for (int k = 0; k < kGroup; k++)
							b[k] = row(b[k].x, b[k].y, (w >> (8 * k)) & 0xffu);
						if (!last) {
							if (to_g) {
#pragma unroll
								for (int k = 0; k < kGroup; k++)
									gfr[(r0 + k) * 32] = b[k];
							} else {
#pragma unroll
								for (int k = 0; k < kGroup; k++)
									xout[k * 32] = b[k];
							}
						}
					} else {
						// ---- 1..3 tail rows of this query ----
#pragma unroll 1
						for (int k = 0; r0 + k < Lq; k++) {
							uint2 b;
							if (first) {
								b = make_uint2(bl, NEG);
								bl += dstep;
							} else {
								b = from_g ? gfr[(r0 + k) * 32] : xin[k * 32];
							}
							b = row(b.x, b.y, (w >> (8 * k)) & 0xffu);
							if (!last) {
								if (to_g)
									gfr[(r0 + k) * 32] = b;
								else
									xout[k * 32] = b;
							}
						}
					}
					// ---- my last row of this strip: results ------------------------
					if (Lq <= r0 + kGroup) {
						int sc0 = 0, sc1 = 0;
						bool have0 = false, have1 = false;
						if (METHOD == M_SW) {
							uint32_t acc = sm.sbest[d_ps][lane];
							acc = O::max3(acc, best, best);
							if (last) {
								sc0 = O::lo(acc);
								sc1 = O::hi(acc);
								have0 = true;
								have1 = PACKED;
							} else {
								sm.sbest[d_ps][lane] = acc;
							}
						} else {
							const int cap0 = d_l0 - 1 - d_c0, cap1 = d_l1 - 1 - d_c0;
							if (cap0 >= 0 && cap0 < K) {
								uint32_t v = 0;
#pragma unroll
								for (int c = 0; c < K; c++)
									if (c == cap0)
										v = Mo[c];
								sc0 = O::lo(v) - gA;
								have0 = true;
							}
							if (PACKED && cap1 >= 0 && cap1 < K) {
								uint32_t v = 0;
#pragma unroll
								for (int c = 0; c < K; c++)
									if (c == cap1)
										v = Mo[c];
								sc1 = O::hi(v) - gA;
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
			}
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
