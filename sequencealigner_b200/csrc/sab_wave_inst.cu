// sab_wave_inst.cu -- instantiations of the strip kernel (sab_wave.cuh) for ONE register-strip
// width, -DSAB_K=<columns>.  build.py compiles it once per width (in parallel: the unrolled kernel
// is ~64 KB of code per instantiation) straight to a cubin, with -DSAB_LG_IMM=<placeholder>; the
// cubins are embedded in libseqalign_b200.so and loaded through sab_module.cu, which writes the
// job's gap constant over the placeholder.  A job only ever loads the widths its subjects'
// lengths ask for (sab_capi.cu: choose_widths).
//
// Packed int16x2 kernels exist for every width; the int32 instantiation (strips that fail the
// range test, rare) only for the widest and the narrowest.
#include "sab_wave.cuh"

#ifndef SAB_K
#error "compile with -DSAB_K=<strip width>"
#endif

namespace sab {

constexpr int kK = SAB_K;
constexpr int kMinb = wave_minb(kK);

template __global__ void sab_wave_kernel<M_NW, true, kK, kMinb>(const TileParams);
template __global__ void sab_wave_kernel<M_GA, true, kK, kMinb>(const TileParams);
template __global__ void sab_wave_kernel<M_SW, true, kK, kMinb>(const TileParams);
#if SAB_K == 32 || SAB_K == 16
template __global__ void sab_wave_kernel<M_NW, false, kK, kMinb>(const TileParams);
template __global__ void sab_wave_kernel<M_GA, false, kK, kMinb>(const TileParams);
template __global__ void sab_wave_kernel<M_SW, false, kK, kMinb>(const TileParams);
#endif

} // namespace sab
