// sab_module.h -- the strip kernels as specialised device images (sab_module.cu)
#pragma once
#include "sab_internal.h"
#include "sab_kernels.cuh"

#include <string>

// ELF64 cubin in img: writes `value` over the placeholder immediate in .text.<kernel>.
// Returns the number of instruction words changed, -1 when the image is not as expected.
int sab_module_patch(unsigned char *img, size_t size, const char *kernel, uint32_t placeholder, uint32_t value);
// embedded device image of register-strip width k (nullptr: no such width)
const unsigned char *sab_module_image(int k, size_t *size);
// mangled name of sab::sab_wave_kernel<method, packed, k, wave_minb(k)>
void sab_module_kernel_name(int method, bool packed, int k, char *out, size_t cap);
// kernel <method, packed, k> with the gap constant `lg` as immediate, loaded into the current
// device's primary context; *cache is per sab_ctx (sab_module_release frees it)
int sab_module_get(void **cache, int method, bool packed, int k, uint32_t lg, int smem, void **fn_out,
		   int *max_blocks_per_sm, std::string *err);
int sab_module_launch(void *fn, int grid, int smem, cudaStream_t stream, const sab::TileParams *p, std::string *err);
void sab_module_release(void *cache);
