// sab_wave_inst.cu -- instantiations of the strip kernel (sab_wave.cuh) for ONE register-strip
// width, -DSAB_K=<columns>.  Compiled once per width (build.py does them in parallel): the
// unrolled kernel is ~64 KB of code per instantiation, and a job only ever launches the
// widths that fit its subjects' lengths (sab_capi.cu: choose_widths).
//
// Packed int16x2 kernels exist for every width; the int32 instantiation (strips that fail the
// range test, rare) only for the widest and the narrowest.
#include "sab_wave.cuh"

#ifndef SAB_K
#error "compile with -DSAB_K=<strip width>"
#endif

namespace sab {

#define SAB_CAT2(a, b) a##b
#define SAB_CAT(a, b) SAB_CAT2(a, b)

constexpr int kK = SAB_K;
// resident blocks per SM the register budget allows: 2K registers hold the strip itself
constexpr int kMinb = kK <= 20 ? 5 : 4;

wave_kernel_t SAB_CAT(wave_kernel_k, SAB_K)(int method, bool packed, int *smem_bytes, int *minb)
{
	*smem_bytes = (int)sizeof(WaveSmem<kK>);
	*minb = kMinb;
	if (packed) {
		switch (method) {
		case M_NW: return sab_wave_kernel<M_NW, true, kK, kMinb>;
		case M_GA: return sab_wave_kernel<M_GA, true, kK, kMinb>;
		default: return sab_wave_kernel<M_SW, true, kK, kMinb>;
		}
	}
#if SAB_K == 32 || SAB_K == 16
	switch (method) {
	case M_NW: return sab_wave_kernel<M_NW, false, kK, kMinb>;
	case M_GA: return sab_wave_kernel<M_GA, false, kK, kMinb>;
	default: return sab_wave_kernel<M_SW, false, kK, kMinb>;
	}
#else
	return nullptr;
#endif
}

} // namespace sab
