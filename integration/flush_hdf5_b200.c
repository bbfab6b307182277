/*
 * flush_hdf5_b200.c -- stands in for src/io/format/hdf5.c in the drop-in build when libhdf5
 * is not available: same three symbols that TU provides to the rest of the reference host
 * (`COMPRESSION`, hdf5.c:12; `FLUSH_REGISTER(FLUSH_HDF5, ...)`, hdf5.c:203; the `-z` flag,
 * hdf5.c:219-232), same output file (/sequences + /similarity_matrix, chunking and deflate
 * rule of hdf5.c:68-89, triangular input expanded to the full matrix), written by the
 * library's own HDF5 writer (sab_h5_write, csrc/sab_h5.cpp) instead of libhdf5.
 */
#include <args.h>
#include <print.h>

#include "io/output.h"
#include "system/os.h"
#include "util/macros.h"

#include "seqalign_b200.h"

uint COMPRESSION;

static void on_chunk_row(int64_t done, int64_t total, void *)
{
	pproportc(done / total, "Converting to HDF5");
}

static bool flush_h5(const struct output *out, const char *path)
{
	pinfo("Writing %zu sequences to HDF5", out->dim);
	if (out->triangular) {
		pinfo("Writing triangular Similarity Matrix to HDF5");
		ppercent(0, "Converting to HDF5");
	} else {
		pinfo("Writing Similarity Matrix to HDF5");
	}
	const int chunk = sab_h5_chunk_dim(out->dim);
	if (chunk)
		pverb("HDF5 chunk size: %d x %d", chunk, chunk);
	const int rc = sab_h5_write(path, out->seqs, out->dim, out->matrix, out->triangular, COMPRESSION,
				    out->triangular ? on_chunk_row : nullptr, nullptr);
	if (rc != SAB_OK) {
		perr("Failed to write HDF5 file: %s", file_name(path));
		return false;
	}
	if (out->triangular)
		ppercent(100, "Converting to HDF5");
	return true;
}
FLUSH_REGISTER(FLUSH_HDF5, flush_h5)

ARG_PARSE_UL(parse_level, 10, uint, (uint), val > 9, "Compression level must be between 0-9")

static void print_level(void)
{
	pinfom("Compression: %u", COMPRESSION);
}

ARG_EXTERN(filter_threshold);
ARG_EXTERN(output_path);

ARGUMENT(compression) = {
	.opt = 'z',
	.lopt = "compression",
	.help = "Compression level for HDF5 datasets [0-9]",
	.param = "N",
	.param_req = ARG_PARAM_REQUIRED,
	.dest = &COMPRESSION,
	.parse_callback = parse_level,
	.action_callback = print_level,
	.action_phase = ARG_CALLBACK_IF_SET,
	.action_order = ARG_ORDER_AFTER(ARG(filter_threshold)),
	.help_order = ARG_ORDER_AFTER(ARG(filter_threshold)),
	ARG_DEPENDS(ARG_RELATION_PARSE, ARG(output_path)),
};
