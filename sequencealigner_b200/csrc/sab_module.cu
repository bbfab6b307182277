// sab_module.cpp -- loads the strip kernels (sab_wave.cuh) as SPECIALISED device images.
//
// The two add-max ops of every DP cell add the same per-lane gap constant (GA: o - e, SW: e).
// As a register operand it costs a register-file read per DPX instruction next to a saturated
// alu pipe; as an immediate operand it is free (measured on B200, profiles/r2_ab_call1.txt:
// C2 +3.2 %, C3 +5.4 %).  The value is a run-time CLI argument (-s / -e), so the kernels are
// compiled once, at build time, with a PLACEHOLDER immediate (SAB_LG_IMM, build.py passes
// -DSAB_LG_IMM=kLgPlaceholder) into one cubin per register-strip width; the cubins are embedded
// in the library (.incbin), and when a context first needs a kernel the placeholder is
// overwritten in the instruction words of THAT kernel's .text section with the job's value and
// the image is handed to the driver (cuModuleLoadData).  sm_100a encodes a 32-bit immediate in
// bits 32..63 of the 128-bit instruction word; every occurrence of the placeholder in the built
// images is such an operand of a VIADDMNMX (tests/test_host.py checks this with cuobjdump).
//
// No compiler runs at load time, nothing is cached on disk, and there is no second code path:
// without the driver entry points the context reports SAB_ERR_CUDA.
#include "sab_module.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#define SAB_EMBED_DECL(K) extern "C" const unsigned char sab_cubin_k##K[], sab_cubin_k##K##_end[];
SAB_WIDTHS(SAB_EMBED_DECL)
#undef SAB_EMBED_DECL

namespace {

struct Image {
	int k;
	const unsigned char *begin, *end;
};
#define SAB_EMBED_ENTRY(K) { K, sab_cubin_k##K, sab_cubin_k##K##_end },
const Image kImages[] = { SAB_WIDTHS(SAB_EMBED_ENTRY) };
#undef SAB_EMBED_ENTRY

// driver entry points, resolved through the runtime (no link-time dependency on libcuda)
struct Driver {
	bool ok = false;
	std::string err;
	decltype(&cuModuleLoadData) ModuleLoadData = nullptr;
	decltype(&cuModuleGetFunction) ModuleGetFunction = nullptr;
	decltype(&cuModuleUnload) ModuleUnload = nullptr;
	decltype(&cuFuncSetAttribute) FuncSetAttribute = nullptr;
	decltype(&cuOccupancyMaxActiveBlocksPerMultiprocessor) Occupancy = nullptr;
	decltype(&cuLaunchKernel) LaunchKernel = nullptr;
	decltype(&cuGetErrorString) GetErrorString = nullptr;
};

template <typename F> bool entry(Driver &d, const char *name, F *out)
{
	void *p = nullptr;
	cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
	const cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
	if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
		cudaGetLastError();
		d.err = std::string("driver entry point ") + name + " not available";
		return false;
	}
	*out = reinterpret_cast<F>(p);
	return true;
}

Driver &driver()
{
	static Driver d;
	static std::once_flag once;
	std::call_once(once, []() {
		d.ok = entry(d, "cuModuleLoadData", &d.ModuleLoadData) &&
		       entry(d, "cuModuleGetFunction", &d.ModuleGetFunction) &&
		       entry(d, "cuModuleUnload", &d.ModuleUnload) &&
		       entry(d, "cuFuncSetAttribute", &d.FuncSetAttribute) &&
		       entry(d, "cuOccupancyMaxActiveBlocksPerMultiprocessor", &d.Occupancy) &&
		       entry(d, "cuLaunchKernel", &d.LaunchKernel) &&
		       entry(d, "cuGetErrorString", &d.GetErrorString);
	});
	return d;
}

std::string cu_error(CUresult r)
{
	const char *s = nullptr;
	Driver &d = driver();
	if (d.GetErrorString && d.GetErrorString(r, &s) == CUDA_SUCCESS && s)
		return s;
	return "CUresult " + std::to_string((int)r);
}

template <typename T> T rd(const unsigned char *p)
{
	T v;
	memcpy(&v, p, sizeof v);
	return v;
}

struct Loaded {
	CUmodule mod = nullptr;
	CUfunction fn = nullptr;
	int max_blocks = 0;
};

struct Cache {
	std::map<long long, Loaded> byKey; // (width index, packed)
};

} // namespace

// Overwrites the placeholder immediate in the .text section of `kernel` (ELF64 cubin in `img`).
// Returns the number of instruction words changed, or -1 when the image is not what this file
// expects (no such section, or the placeholder somewhere that is not an immediate field).
int sab_module_patch(unsigned char *img, size_t size, const char *kernel, uint32_t placeholder, uint32_t value)
{
	if (size < 0x40 || memcmp(img, "\177ELF", 4) != 0 || img[4] != 2 /* ELFCLASS64 */)
		return -1;
	const uint64_t shoff = rd<uint64_t>(img + 0x28);
	const unsigned shentsize = rd<uint16_t>(img + 0x3a), shnum = rd<uint16_t>(img + 0x3c),
		       shstrndx = rd<uint16_t>(img + 0x3e);
	if (!shoff || shentsize < 0x40 || shstrndx >= shnum || shoff + (uint64_t)shentsize * shnum > size)
		return -1;
	auto sh = [&](unsigned i) { return img + shoff + (uint64_t)shentsize * i; };
	const uint64_t str_off = rd<uint64_t>(sh(shstrndx) + 0x18), str_size = rd<uint64_t>(sh(shstrndx) + 0x20);
	if (str_off + str_size > size)
		return -1;
	const std::string want = std::string(".text.") + kernel;
	for (unsigned i = 0; i < shnum; i++) {
		const uint32_t name = rd<uint32_t>(sh(i));
		if (name >= str_size || want != reinterpret_cast<const char *>(img + str_off + name))
			continue;
		const uint64_t off = rd<uint64_t>(sh(i) + 0x18), len = rd<uint64_t>(sh(i) + 0x20);
		if (off + len > size || (len & 15))
			return -1;
		int n = 0;
		for (uint64_t o = 0; o + 16 <= len; o += 16) {
			unsigned char *ins = img + off + o;
			for (int b = 0; b < 16; b += 4) {
				if (rd<uint32_t>(ins + b) != placeholder)
					continue;
				if (b != 4)
					return -1; // not the immediate field of an instruction word
				memcpy(ins + 4, &value, 4);
				n++;
			}
		}
		return n;
	}
	return -1;
}

const unsigned char *sab_module_image(int k, size_t *size)
{
	for (const Image &im : kImages)
		if (im.k == k) {
			*size = (size_t)(im.end - im.begin);
			return im.begin;
		}
	return nullptr;
}

void sab_module_kernel_name(int method, bool packed, int k, char *out, size_t cap)
{
	// sab::sab_wave_kernel<METHOD, PACKED, K, MINB>(sab::TileParams), Itanium mangling
	snprintf(out, cap, "_ZN3sab15sab_wave_kernelILi%dELb%dELi%dELi%dEEEvNS_10TileParamsE", method, packed ? 1 : 0, k,
		 sab::wave_minb(k));
}

// The kernel <method, packed, width k> specialised for `lg`, loaded into the CURRENT context (the
// caller has made its device current).  `*cache` belongs to one sab_ctx.
int sab_module_get(void **cache, int method, bool packed, int k, uint32_t lg, int smem, void **fn_out,
		   int *max_blocks_per_sm, std::string *err)
{
	Driver &d = driver();
	if (!d.ok) {
		*err = d.err;
		return SAB_ERR_CUDA;
	}
	if (!*cache)
		*cache = new Cache;
	Cache &c = *static_cast<Cache *>(*cache);
	const long long key = (long long)k * 2 + (packed ? 1 : 0);
	auto it = c.byKey.find(key);
	if (it == c.byKey.end()) {
		size_t size = 0;
		const unsigned char *src = sab_module_image(k, &size);
		if (!src) {
			*err = "no device image of width " + std::to_string(k);
			return SAB_ERR_STATE;
		}
		char name[160];
		sab_module_kernel_name(method, packed, k, name, sizeof name);
		std::vector<unsigned char> img(src, src + size);
		const int n = sab_module_patch(img.data(), img.size(), name, sab::kLgPlaceholder, lg);
		if (n < 0 || (n == 0 && method != sab::M_NW)) {
			*err = std::string("device image of ") + name + " could not be specialised";
			return SAB_ERR_STATE;
		}
		Loaded l;
		CUresult r = d.ModuleLoadData(&l.mod, img.data());
		if (r == CUDA_SUCCESS)
			r = d.ModuleGetFunction(&l.fn, l.mod, name);
		if (r == CUDA_SUCCESS)
			r = d.FuncSetAttribute(l.fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, smem);
		if (r == CUDA_SUCCESS)
			r = d.Occupancy(&l.max_blocks, l.fn, sab::kWaveThreads, (size_t)smem);
		if (r != CUDA_SUCCESS) {
			*err = std::string(name) + ": " + cu_error(r);
			if (l.mod)
				d.ModuleUnload(l.mod);
			return r == CUDA_ERROR_OUT_OF_MEMORY ? SAB_ERR_NOMEM : SAB_ERR_CUDA;
		}
		it = c.byKey.emplace(key, l).first;
	}
	*fn_out = it->second.fn;
	*max_blocks_per_sm = it->second.max_blocks;
	return SAB_OK;
}

int sab_module_launch(void *fn, int grid, int smem, cudaStream_t stream, const sab::TileParams *p, std::string *err)
{
	Driver &d = driver();
	void *args[] = { const_cast<sab::TileParams *>(p) };
	const CUresult r = d.LaunchKernel(static_cast<CUfunction>(fn), (unsigned)grid, 1, 1, sab::kWaveThreads, 1, 1,
					  (unsigned)smem, reinterpret_cast<CUstream>(stream), args, nullptr);
	if (r != CUDA_SUCCESS) {
		*err = "cuLaunchKernel: " + cu_error(r);
		return SAB_ERR_CUDA;
	}
	return SAB_OK;
}

// The owning context's device must be current (modules live in its primary context).
void sab_module_release(void *cache)
{
	if (!cache)
		return;
	Cache *c = static_cast<Cache *>(cache);
	Driver &d = driver();
	for (auto &kv : c->byKey)
		if (kv.second.mod && d.ModuleUnload)
			d.ModuleUnload(kv.second.mod);
	delete c;
}

// Introspection (include/seqalign_b200.h): the specialised image a context would load.
extern "C" int sab_kernel_image(int method, int packed, int width, uint32_t gap_constant, unsigned char *out,
				size_t cap, size_t *size, char *kernel_name, size_t name_cap)
{
	size_t n = 0;
	const unsigned char *src = sab_module_image(width, &n);
	if (!src || method < sab::M_NW || method > sab::M_SW || (!packed && width != 32 && width != 16))
		return SAB_ERR_ARG;
	if (size)
		*size = n;
	char name[160];
	sab_module_kernel_name(method, packed != 0, width, name, sizeof name);
	if (kernel_name && name_cap)
		snprintf(kernel_name, name_cap, "%s", name);
	if (!out)
		return 0;
	if (cap < n)
		return SAB_ERR_ARG;
	memcpy(out, src, n);
	const int sites = sab_module_patch(out, n, name, sab::kLgPlaceholder, gap_constant);
	return sites < 0 ? SAB_ERR_STATE : sites;
}
