/*
 * source_b200.c -- stands in for src/io/source/fasta.c and src/io/source/dsv.c in the drop-in
 * build: the two parser plugins of the reference's `sources` registry (src/io/source.h:14-23,
 * walked by input_load, src/io/input.c:47-57) with the text processing done by the library
 * (sab_parse_fasta / sab_parse_dsv, csrc/sab_source.cpp: one table-driven pass, in place).
 * Accepted inputs, the resulting store and every error line are the reference's
 * (tests/test_sources.py, tests/test_dropin.py).
 *
 * What stays here is what only the host can do: talk to the user.  A delimiter-separated file
 * whose first row names no sequence column makes the reference ask which column to read and
 * whether that row is a header at all (dsv.c:146-160); the same two prompts are issued here,
 * between sab_dsv_header and sab_parse_dsv.
 */
#include <print.h>
#include <stdlib.h>
#include <string.h>

#include "bio/align.h"
#include "io/input.h"
#include "io/source.h"
#include "system/os.h"
#include "util/benchmark.h"

#include "seqalign_b200.h"

static enum parse_result finish(int rc, const char *msg, struct input *in, s32 num, s32 longest,
				const char *what)
{
	if (rc == SAB_ERR_PARSE) {
		perr("%s", msg);
		return PARSER_ERROR;
	}
	if (rc != SAB_OK) {
		perr("%s input rejected by libseqalign_b200 (%d)", what, rc);
		return PARSER_ERROR;
	}
	in->num = num;
	in->max = longest;
	pverbl("%s parsing finished successfuly", what);
	return PARSER_SUCCESS;
}

static enum parse_result fasta_b200(struct source src, struct input *in)
{
	pverbm("Trying out FASTA parser");
	if (sab_source_format(src.ext, nullptr) != SAB_SRC_FASTA)
		return PARSER_UNSUPPORTED;
	pverbm("Using FASTA parser");
	char msg[128] = "";
	s32 num = 0, longest = 0;
	static_assert(sizeof(SEQ_LUT[0]) == sizeof(int32_t));
	const int rc = sab_parse_fasta(src.file, (size_t)(src.fend - src.file), (const int32_t *)SEQ_LUT,
				       GAP_PEN, &num, &longest, msg, sizeof(msg));
	return finish(rc, msg, in, num, longest, "FASTA");
}

static enum parse_result dsv_b200(struct source src, struct input *in)
{
	pverbm("Trying out DSV parser");
	uint8_t delim = 0;
	if (sab_source_format(src.ext, &delim) != SAB_SRC_DSV)
		return PARSER_UNSUPPORTED;
	pverbm("Using DSV parser");

	const size_t bytes = (size_t)(src.fend - src.file);
	char msg[128] = "";
	s32 cols = 0, seq_col = -1;
	int rc = sab_dsv_header(src.file, bytes, delim, &cols, &seq_col, nullptr, 0, msg, sizeof(msg));
	if (rc != SAB_OK)
		return finish(rc, msg, in, 0, 0, "DSV");

	int first_row_is_data = 0;
	if (seq_col < 0) {
		/* nobody named the column: ask, outside the timed input phase (dsv.c:146-160).  Only now are the
		 * first row's field texts needed (they cannot be longer than the file) */
		char *names = malloc(bytes + 1);
		const char **list = malloc(sizeof(*list) * ((size_t)cols + 1));
		if (!names || !list ||
		    sab_dsv_header(src.file, bytes, delim, &cols, &seq_col, names, bytes + 1, msg, sizeof(msg)) != SAB_OK) {
			free(names);
			free(list);
			perr("Out of memory during DSV parsing");
			return PARSER_ERROR;
		}
		const char *name = names;
		for (s32 c = 0; c < cols; c++, name += strlen(name) + 1)
			list[c] = name;
		list[cols] = "No header line";
		bench_input_end();
		pinfol("Which column contains your sequences?");
		s32 pick = pchoice(list, cols + 1, "Enter column number");
		if (pick == cols) {
			first_row_is_data = 1;
			/* a one-entry list cannot be asked about (clix refuses it): it is that column */
			pick = 0;
			if (cols > 1) {
				pinfol("Which column contains a sequence?");
				pick = pchoice(list, cols, "Enter column number");
			}
		}
		bench_input_start();
		free(list);
		free(names);
		seq_col = pick;
	}

	s32 num = 0, longest = 0;
	rc = sab_parse_dsv(src.file, bytes, delim, seq_col, first_row_is_data, (const int32_t *)SEQ_LUT, GAP_PEN,
			   &num, &longest, msg, sizeof(msg));
	return finish(rc, msg, in, num, longest, "DSV");
}

SOURCE_REGISTER(fasta_b200, fasta_b200);
SOURCE_REGISTER(dsv_b200, dsv_b200);
