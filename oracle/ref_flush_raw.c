/*
 * TEST INFRASTRUCTURE ONLY (oracle/): never linked into the product library.
 *
 * Stand-in output writer used when the UNMODIFIED reference CPU sources under
 * /root/reference/src are compiled into oracle/_ref/seqalign.  The reference's
 * own writer (src/io/format/hdf5.c) needs libhdf5, which this image does not
 * have; every other reference translation unit is compiled as-is.  This TU
 * provides exactly the three symbols the other reference TUs expect from the
 * writer TU (reference: src/io/format/hdf5.c:12 `COMPRESSION`,
 * hdf5.c:203 `FLUSH_REGISTER(FLUSH_HDF5, ...)`, hdf5.c:219 `ARGUMENT(compression)`)
 * and dumps the similarity matrix raw so tests can compare it bit for bit:
 *
 *   u64 magic "SQALRAW1" | u64 dim | u64 triangular |
 *   s32 matrix[dim*dim or dim*(dim-1)/2] |
 *   dim NUL-terminated sequences (the post-filter /sequences list)
 */
#include <args.h>
#include <print.h>
#include <stdio.h>
#include <string.h>

#include "io/output.h"
#include "system/os.h"
#include "util/macros.h"

uint COMPRESSION;

static bool flush_raw(const struct output *out, const char *path)
{
	FILE *f = fopen(path, "wb");
	if (!f) {
		perr("Failed to create raw dump: %s", file_name(path));
		return false;
	}
	unsigned long long hdr[3] = { 0x31574152'4C415153ULL,
				      (unsigned long long)out->dim,
				      (unsigned long long)out->triangular };
	size_t n = out->triangular ? alignments(out->dim) : out->dim * out->dim;
	bool ok = fwrite(hdr, sizeof(hdr), 1, f) == 1;
	ok = ok && (n == 0 || fwrite(out->matrix, sizeof(s32), n, f) == n);
	for (size_t i = 0; ok && i < out->dim; i++) {
		size_t l = strlen(out->seqs[i]) + 1;
		ok = fwrite(out->seqs[i], 1, l, f) == l;
	}
	ok = (fclose(f) == 0) && ok;
	if (!ok)
		perr("Failed to write raw dump: %s", file_name(path));
	return ok;
}
FLUSH_REGISTER(FLUSH_HDF5, flush_raw)

ARG_PARSE_UL(parse_level, 10, uint, (uint), val > 9,
	     "Compression level must be between 0-9")

ARG_EXTERN(filter_threshold);
ARG_EXTERN(output_path);

ARGUMENT(compression) = {
	.opt = 'z',
	.lopt = "compression",
	.help = "Compression level (accepted, ignored by the raw dump)",
	.param = "N",
	.param_req = ARG_PARAM_REQUIRED,
	.dest = &COMPRESSION,
	.parse_callback = parse_level,
	.action_order = ARG_ORDER_AFTER(ARG(filter_threshold)),
	.help_order = ARG_ORDER_AFTER(ARG(filter_threshold)),
	ARG_DEPENDS(ARG_RELATION_PARSE, ARG(output_path)),
};
