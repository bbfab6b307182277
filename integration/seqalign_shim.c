/*
 * seqalign_shim.c -- the reference-side binding of libseqalign_b200.so.
 *
 * Drop-in replacement for ONE translation unit of jakovdev/SequenceAligner,
 * src/interface/seqalign_cuda.c (and with it src/bio/kernels.cu): it exports the
 * two functions the reference host calls, with the reference's exact signatures
 * (src/interface/seqalign_cuda.h:7-9)
 *
 *     bool cuda_memory(size_t bytes);                 <- output_load, src/io/output.c:37
 *     bool cuda_align(struct input, struct output);   <- main,        src/main.c:31
 *
 * plus the `-C, --no-cuda` flag that TU owns (seqalign_cuda.c:307-316) and the three
 * `kernel_*` symbols the method plugins name in `.kernel` when built with USE_CUDA
 * (src/bio/align.h:47-51, src/bio/method/{nw,ga,sw}.c).  Everything else of the host
 * -- CLI, parsers, filter, output_load, HDF5 writer, the `struct align` plugin
 * registry -- is the unmodified reference code (integration/Makefile compiles it in
 * place from the reference tree with -DUSE_CUDA).
 *
 * What crosses into the library is only the C ABI of include/seqalign_b200.h.
 */
#ifdef _OPENMP
#include <omp.h>
#endif
#include "interface/seqalign_cuda.h"

#include <args.h>
#include <print.h>
#include <stdlib.h>
#include <string.h>

#include "bio/align.h"
#include "io/input.h"
#include "io/output.h"
#include "system/os.h"
#include "util/benchmark.h"
#include "util/macros.h"

#include "seqalign_b200.h"

[[gnu::nonnull]]
bool align(struct input, struct output); /* the reference's CPU path, src/bio/align.c:21 */

/*
 * The method plugins identify their device kernel by symbol (`.kernel = kernel_nw`).
 * The B200 library selects by method id instead, so these are address tags only:
 * they are never launched or called.
 */
void kernel_nw(s32 *, s64, s64)
{
}
void kernel_ga(s32 *, s64, s64)
{
}
void kernel_sw(s32 *, s64, s64)
{
}

static bool no_cuda; /* -C: the user asked for the reference's CPU path */
static int device_count = -1;
static sab_multi *multi; /* one set of device contexts for the filter and the alignment */

bool b200_cuda_disabled(void)
{
	return no_cuda;
}

static bool device_probe(void)
{
	if (device_count < 0) {
		device_count = sab_device_count();
		if (device_count > 0)
			pverb("Available B200 devices: %d", device_count);
	}
	if (device_count == 0) {
		/* no silent fallback: the user can pass -C to run the reference's CPU code */
		perr("No sm_100 (B200) CUDA device available");
		return false;
	}
	return true;
}

/*
 * seqalign_cuda.c:71-93 answers "does the full matrix (x 4/3) fit in VRAM?" because the
 * reference keeps the whole score matrix on the device.  This backend streams row strips
 * and never holds more than two of them, so the device never forces the triangular
 * layout; the host-RAM rule in output_load (src/io/output.c:36) alone decides.
 */
bool cuda_memory(size_t bytes)
{
	if (no_cuda)
		return true;
	if (!device_probe())
		exit(EXIT_FAILURE);
	size_t free_b = 0, total_b = 0;
	if (sab_device_memory(sab_device_ordinal(0), &free_b, &total_b) != SAB_OK) {
		perr("Failed to query device memory");
		exit(EXIT_FAILURE);
	}
	pverb("GPU: %.2f GiB free / %.2f GiB total (matrix %.2f GiB stays on the host)",
	      (double)free_b / (double)GiB, (double)total_b / (double)GiB,
	      (double)bytes / (double)GiB);
	return true;
}

static int method_id(void)
{
	if (ALIGN->kernel == (const void *)kernel_nw)
		return SAB_NW;
	if (ALIGN->kernel == (const void *)kernel_ga)
		return SAB_GA;
	if (ALIGN->kernel == (const void *)kernel_sw)
		return SAB_SW;
	return -1;
}

static void on_progress(int64_t done, int64_t total, void *)
{
	pproportc(done / total, "Aligning sequences");
}

static void context_release(void)
{
	sab_multi_destroy(multi);
	multi = nullptr;
}

/*
 * The B200s of the box behind one handle: a context + host thread per device, created lazily
 * by the library and only as many as the job can keep busy.  Shared by filter() (which runs
 * first, inside input_load) and cuda_align().  All arguments are parsed by then, so ALIGN,
 * SEQ_LUT, SUB_MAT and the gap values are final.
 */
sab_multi *b200_context(void)
{
	if (multi)
		return multi;
	if (!device_probe())
		return nullptr;
	sab_config cfg = {
		.method = method_id(),
		/* validate_ga parks the unused values at SCORE_MIN when it switches to NW (ga.c:83-84) */
		.gap_pen = GAP_PEN == SCORE_MIN ? 0 : GAP_PEN,
		.gap_open = GAP_OPN == SCORE_MIN ? 0 : GAP_OPN,
		.gap_ext = GAP_EXT == SCORE_MIN ? 0 : GAP_EXT,
#ifdef _OPENMP
		/* -T (os.c:466-473 has already applied it to the OpenMP runtime) also bounds the library's
		 * host threads (sequence packing, copies into out.matrix) */
		.host_threads = omp_get_max_threads(),
#endif
	};
	if (cfg.method < 0) {
		perr("Alignment method has no B200 kernel");
		return nullptr;
	}
	static_assert(sizeof(cfg.seq_lut) == sizeof(SEQ_LUT));
	static_assert(sizeof(cfg.sub_mat) == sizeof(SUB_MAT));
	static_assert(sizeof(sab_meta) == sizeof(struct meta));
	memcpy(cfg.seq_lut, SEQ_LUT, sizeof(SEQ_LUT));
	memcpy(cfg.sub_mat, SUB_MAT, sizeof(SUB_MAT));
	if (sab_multi_create(&multi, &cfg, 0) != SAB_OK) {
		perr("%s", sab_multi_last_error(multi));
		context_release();
		return nullptr;
	}
	atexit(context_release);
	return multi;
}

bool cuda_align(struct input in, struct output out)
{
	if (no_cuda)
		return align(in, out);
	sab_multi *m = b200_context();
	if (!m)
		return false;
	bool ok = false;
	if (sab_multi_upload(m, in.seqs, (const sab_meta *)in.meta, in.num, in.max) != SAB_OK) {
		perr("%s", sab_multi_last_error(m));
		goto done;
	}
	pinfo("Using %d B200 device%s", sab_multi_device_count(m),
	      sab_multi_device_count(m) == 1 ? "" : "s");

	pinfo("Performing %zu pairwise alignments", (size_t)alignments((s64)in.num));
	ppercent(0, "Aligning sequences");
	bench_align_start();
	/* out.matrix is NULL under -W (src/io/output.c:18): compute, copy nothing back */
	const int rc = sab_multi_align(m, out.matrix, out.triangular, on_progress, nullptr);
	bench_align_end();
	if (rc != SAB_OK) {
		perr("%s", sab_multi_last_error(m));
		goto done;
	}
	ppercent(100, "Aligning sequences");
	bench_align_print();
	{
		sab_stats st;
		if (sab_multi_get_stats(m, &st) == SAB_OK)
			pverb("B200: %.1f GCUPS kernel, %lld kernel launches, %lld int32 re-runs",
			      (double)st.cells / (st.kernel_ms * 1e6), (long long)st.launches,
			      (long long)st.rerun_pairs);
	}
	ok = true;
done:
	return ok; /* the contexts are released at exit (context_release) */
}

static void print_backend(void)
{
	pinfom("CUDA: Enabled (libseqalign_b200, sm_100a)");
}

ARG_EXTERN(compression);
ARG_EXTERN(threads);

ARGUMENT(disable_cuda) = {
	.opt = 'C',
	.lopt = "no-cuda",
	.help = "Disable CUDA",
	.set = &no_cuda,
	.action_callback = print_backend,
	.action_phase = ARG_CALLBACK_IF_UNSET,
	.action_order = ARG_ORDER_AFTER(ARG(compression)),
	.help_order = ARG_ORDER_AFTER(ARG(threads)),
};
