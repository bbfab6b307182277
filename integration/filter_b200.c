/*
 * filter_b200.c -- stands in for src/bio/filter.c in the drop-in build: `filter()` (called by
 * input_load, src/io/input.c:84) takes its decisions from the library (sab_multi_filter: the
 * letter compares of filter.c:38-55 on the B200s, the order-dependent part resolved
 * sequentially) and then compacts the sequence store exactly as filter.c:64-77 does.
 *
 * The reference's own translation unit is compiled INTO this one, in place from the reference
 * tree and with its entry point renamed: it keeps owning the `-f` flag (filter.c:91-117) and its
 * `threshold`, and its CPU filter stays reachable under `-C, --no-cuda` -- the reference CLI's
 * explicit switch to the reference's CPU code, like the shim's cuda_align().  Note that the
 * reference's OpenMP loop races on lost[] (filter.c:20,35-54): with more than one thread its
 * result depends on timing; the library reproduces the sequential (`-T 1`) result.
 */
#define filter filter_reference_cpu
#include "bio/filter.c" /* -I$(REF)/src */
#undef filter

#include "seqalign_b200.h"

/* integration/seqalign_shim.c */
bool b200_cuda_disabled(void);
sab_multi *b200_context(void);

bool filter(struct input *in)
{
	if (threshold <= 0.0f)
		return true;
	if (b200_cuda_disabled())
		return filter_reference_cpu(in);

	const s32 before = in->num;
	unsigned char *drop = calloc((size_t)before, 1);
	if (!drop) {
		perr("Out of memory during sequence filtering");
		return false;
	}
	sab_multi *m = b200_context();
	if (!m) {
		free(drop);
		return false;
	}
	ppercent(0, "Filtering sequences");
	bench_filter_start();
	static_assert(sizeof(sab_meta) == sizeof(struct meta));
	const int rc = sab_multi_filter(m, in->seqs, (const sab_meta *)in->meta, before, threshold, drop);
	if (rc < 0) {
		perr("%s", sab_multi_last_error(m));
		free(drop);
		return false;
	}
	/* survivors slide down over the dropped ones, order kept, NUL terminators included */
	s32 kept = 0, longest = 0, tail = 0;
	for (s32 k = 0; k < before; k++) {
		if (drop[k])
			continue;
		struct meta cur = in->meta[k];
		if (cur.off != tail)
			memmove(in->seqs + tail, in->seqs + cur.off, (size_t)cur.len + 1);
		cur.off = tail;
		tail += cur.len + 1;
		in->meta[kept++] = cur;
		if (cur.len > longest)
			longest = cur.len;
	}
	in->num = kept;
	in->max = longest;
	free(drop);
	bench_filter_end();
	ppercent(100, "Filtering sequences");

	pinfo("Filtered out %d sequences", before - kept);
	if (kept < SEQ_N_MIN) {
		perr("Not enough sequences: %d (min: %d)", kept, SEQ_N_MIN);
		return false;
	}
	bench_filter_print();
	return true;
}
