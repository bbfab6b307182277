/*
 * seqalign_b200.h -- C ABI of libseqalign_b200.so
 *
 * B200-native (sm_100a) replacement for ONE hot path of jakovdev/SequenceAligner:
 * the score-only all-vs-all pairwise DP fill (Needleman-Wunsch linear gap, Gotoh
 * global affine, Smith-Waterman local affine) over every i<j pair of a sequence
 * set.  Plain pointers and sizes only; no C++ / torch types cross this boundary.
 *
 * What each entry point replaces in the reference (paths under the reference
 * repo root):
 *
 *   sab_device_count / sab_device_memory
 *        <- cuda_device_init + cuda_memory   src/interface/seqalign_cuda.c:48-93
 *   sab_create
 *        <- `struct constants C` fill + cudaMemcpyToSymbol
 *                                            src/interface/seqalign_cuda.c:115-125,168
 *           (method selection = ALIGN->kernel / ALIGN->gap, src/bio/align.h:28-38)
 *   sab_upload
 *        <- cudaMalloc/cudaMemcpy of letters + meta
 *                                            src/interface/seqalign_cuda.c:127-132
 *   sab_align / sab_align_shard
 *        <- the batched launch + D2H loop    src/interface/seqalign_cuda.c:134-294
 *           and, semantically, align() + output_fill()
 *                                            src/bio/align.c:21-72, src/io/output.c:68-84
 *   sab_filter / sab_filter_matches + sab_filter_resolve / sab_multi_filter
 *        <- filter() decisions               src/bio/filter.c:14-62 (sequential -T 1 semantics)
 *   sab_matrix_lookup
 *        <- parse_matrix                     src/bio/matrices.c:44-57
 *   sab_h5_write
 *        <- flush_hdf5                       src/io/format/hdf5.c:14-202
 *   sab_source_format / sab_parse_fasta / sab_dsv_header / sab_parse_dsv / sab_build_meta
 *        <- parse_fasta, parse_dsv           src/io/source/fasta.c:16-85, src/io/source/dsv.c:80-230
 *           and the meta pass of input_load  src/io/input.c:70-81
 *
 * The reference-signature shim on top of this ABI (`bool cuda_align(struct input,
 * struct output)`, `bool cuda_memory(size_t)`, src/interface/seqalign_cuda.h:7-9)
 * is in integration/seqalign_shim.c; INTEGRATION.md shows how a
 * reference maintainer links it.
 *
 * There is NO CPU fallback: every compute entry point fails with
 * SAB_ERR_NODEVICE / SAB_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef SEQALIGN_B200_H
#define SEQALIGN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SAB_ABI_VERSION 3

enum sab_status {
	SAB_OK = 0,
	SAB_ERR_ARG = -1, /* bad argument / unsupported configuration */
	SAB_ERR_NODEVICE = -2, /* no usable CUDA device */
	SAB_ERR_CUDA = -3, /* CUDA runtime error, see sab_last_error */
	SAB_ERR_NOMEM = -4, /* host or device allocation failed */
	SAB_ERR_STATE = -5, /* call order violated (e.g. align before upload) */
	SAB_ERR_OVERFLOW = -6, /* sab_filter_matches only: more matching pairs than the caller's list holds
				* (the alignment path never returns it: an SW strip whose int16 re-run
				* list runs full is swept again in int32) */
	SAB_ERR_PARSE = -7, /* sab_parse_* / sab_dsv_header: the input text is rejected; `err` holds the
			     * reference's own message for it (e.g. "Sequence #3 is invalid") */
};

/* src/bio/method/{nw,ga,sw}.c */
enum sab_method { SAB_NW = 0, SAB_GA = 1, SAB_SW = 2 };

/* == struct meta, src/bio/align.h:6-9 */
typedef struct sab_meta {
	int32_t off; /* byte offset of the sequence in `seqs` */
	int32_t len; /* letters, excluding the terminating NUL */
} sab_meta;

enum sab_flags {
	SAB_FLAG_FORCE_INT32 = 1, /* testing: never use the packed int16x2 kernels */
	SAB_FLAG_FORCE_INT16 = 2, /* testing: skip the a-priori range test (SW still re-runs) */
	SAB_FLAG_STRIP32 = 8, /* testing/tuning: always use the 32-column register-strip kernel */
	SAB_FLAG_STRIP16 = 16, /* testing/tuning: always use the 16-column register-strip kernel */
	SAB_FLAG_ORDERED = 32, /* issue row strips in ascending row order instead of costliest first: the
				* writes into `matrix` then advance through ascending offsets (a file-backed
				* MAP_SHARED sink, src/io/output.c:35-55, gets sequential write-back) */
	SAB_FLAG_TEST_RERUN_CAP = 64, /* testing: SW int16 re-run list of 4 entries (exercises the int32 re-sweep) */
	SAB_FLAG_FILTER_SEQUENTIAL = 128, /* testing/tuning: sab_filter takes the panel-sequential path directly */
	SAB_FLAG_TEST_FILTER_CAP = 256, /* testing: filter match list of 16 entries (exercises the fallback) */
	SAB_FLAG_PER_CHUNK_WIDTHS = 1024, /* tuning: a register-strip width per subject chunk (one launch per width in
					  * use, concurrent) instead of one width per row strip */
	SAB_FLAG_SERIAL_WIDTHS = 512, /* tuning, with PER_CHUNK_WIDTHS: the per-width launches of a row strip run back to
				       * back on one stream instead of concurrently on side streams */
};

typedef struct sab_config {
	int32_t method; /* enum sab_method */
	int32_t gap_pen; /* <= 0, linear (GAP_PEN, src/bio/align.c:14)  */
	int32_t gap_open; /* <= 0, affine (GAP_OPN, src/bio/align.c:15)  */
	int32_t gap_ext; /* <= 0, affine (GAP_EXT, src/bio/align.c:16)  */
	int32_t seq_lut[128]; /* SEQ_LUT, ASCII -> index or -1 (src/bio/matrices.c:12) */
	int32_t sub_mat[24 * 24]; /* SUB_MAT, must be symmetric (src/bio/matrices.c:13) */
	int32_t device; /* CUDA device ordinal this context drives */
	int32_t panel; /* sequences per panel (= rows per row strip), multiple of 128; 0 = auto */
	int32_t flags; /* enum sab_flags */
	int32_t shards_hint; /* how many shards sab_align_shard will be called with (sizes the
			      * auto panel so every shard gets several row strips); 0 = 1 */
	int32_t host_threads; /* host threads for packing the sequences and copying scores into `matrix`
			       * (the reference's -T, src/system/os.c:466-473); 0 = min(cores, 8) */
	int32_t reserved[3];
} sab_config;

typedef struct sab_stats {
	double kernel_ms; /* device time covered by the alignment kernels of the last sab_align*
			   * (union of the per-strip CUDA-event intervals; strips overlap) */
	double busy_ms; /* sum of the per-strip kernel durations */
	double device_ms; /* CUDA-event time from the first kernel's start to the end of all
			   * device work of the call (kernels, re-runs, D2H), gaps included */
	double total_ms; /* wall time of the last sab_align* call */
	double upload_ms; /* wall time of the last sab_upload */
	int64_t pairs; /* pairs scored by the last sab_align* call */
	int64_t cells; /* sum of len_i*len_j over those pairs */
	int64_t rerun_pairs; /* pairs re-scored in int32 after int16 overflow */
	int64_t launches; /* kernels launched by the last sab_align* call */
	int64_t tiles; /* panel tiles processed */
	int64_t tiles_int32; /* panel tiles routed wholesale to the int32 kernel */
	int64_t h2d_bytes; /* bytes copied host->device by the last sab_upload */
	int64_t d2h_bytes; /* bytes copied device->host by the last sab_align* */
	int64_t rerun_strips; /* SW strips swept again in int32 because their re-run list ran full */
	int64_t filter_matches; /* pairs that passed the threshold in the last sab_filter* (-1: sequential path) */
	double filter_ms; /* wall time of the last sab_filter* / sab_multi_filter */
	int32_t width_launches[16]; /* kernel launches per register-strip width, widest (32) first in steps of
				     * 2 columns down to 16: a row strip runs at the width that pads its
				     * subjects least (under SAB_FLAG_PER_CHUNK_WIDTHS: once per width in use) */
} sab_stats;

typedef struct sab_ctx sab_ctx;

/* Called from the thread that called sab_align*, between tiles. */
typedef void (*sab_progress_fn)(int64_t pairs_done, int64_t pairs_total, void *user);

int sab_abi_version(void);

/* Number of usable CUDA devices, i.e. of compute capability 10.x (0 when none / driver missing). */
int sab_device_count(void);
/* CUDA ordinal of the index-th usable device (what sab_config.device takes), or SAB_ERR_NODEVICE. */
int sab_device_ordinal(int index);
/* cudaMemGetInfo of CUDA ordinal `device`. */
int sab_device_memory(int device, size_t *free_bytes, size_t *total_bytes);

int sab_create(sab_ctx **ctx, const sab_config *cfg);
void sab_destroy(sab_ctx *ctx);
const char *sab_last_error(const sab_ctx *ctx);

/*
 * Hands the host sequence store to the device: `seqs` holds `num` uppercase
 * ASCII sequences, NUL-terminated, back to back; meta[k] locates sequence k;
 * `max` is the longest length (== struct input, src/io/input.h:6-11).  The
 * letters are mapped through seq_lut, length-binned (stable sort inside each
 * panel) and packed into the two device layouts of DESIGN.md section 3.  The
 * host buffers are not referenced after return.  May be called again on the
 * same context (a new set replaces the old one; device buffers are re-used
 * when large enough and released by sab_destroy).
 */
int sab_upload(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
	       int32_t max);

/*
 * Scores every pair i<j and fills the caller-owned host matrix exactly like
 * output_fill (src/io/output.c:68-84):
 *   triangular != 0 : matrix[j(j-1)/2 + i] = score(i,j)       (length num(num-1)/2)
 *   triangular == 0 : matrix[num*i+j] = matrix[num*j+i] = score(i,j); the
 *                     diagonal is not written.
 * matrix == NULL computes (and keeps the scores in device memory) but copies
 * nothing back, like the reference's -W / --no-write (seqalign_cuda.c:233).
 * `matrix` may be pageable or a file-backed MAP_SHARED mapping: it is only
 * written with plain host stores from this call's worker threads.
 */
int sab_align(sab_ctx *ctx, int32_t *matrix, int triangular, sab_progress_fn progress,
	      void *user);

/*
 * Multi-GPU: the panel-tile list of the job is split into `nshards`
 * cost-balanced shards; this call computes shard `shard` on this context's
 * device and writes only that shard's cells of `matrix` (disjoint between
 * shards, so several contexts/threads/processes can fill one matrix, or each
 * its own).  sab_align == shard 0 of 1.
 */
int sab_align_shard(sab_ctx *ctx, int32_t shard, int32_t nshards, int32_t *matrix,
		    int triangular, sab_progress_fn progress, void *user);

/* Row strips of the job (fixed by sab_upload): strip k covers original rows
 * [row0, row0 + rows).  sab_strip_owner tells which shard computes it, so a
 * multi-process host knows which rows of `matrix` each rank filled. */
int sab_strip_count(const sab_ctx *ctx);
/* Host only (no device needed): the same plan computed from the lengths alone.
 * panel = 0 sizes the panel for `nshards` shards (== sab_config.shards_hint = nshards).
 * Returns the number of strips; fills at most max_strips entries. */
int sab_plan_strips(const sab_meta *meta, int32_t num, int32_t panel, int32_t nshards,
		    int32_t max_strips, int32_t *row0, int32_t *rows, int32_t *owner);
int sab_strip_info(const sab_ctx *ctx, int32_t strip, int32_t nshards, int32_t *row0,
		   int32_t *rows, int32_t *owner);

int sab_get_stats(const sab_ctx *ctx, sab_stats *out);

/* The strip kernels ship as one sm_100a device image (cubin) per register-strip width, embedded
 * in the library; the per-lane gap constant of the cell update (GA: open - extend, SW: extend;
 * two int16 lanes or one int32) is an IMMEDIATE operand that is written into the instruction
 * words when a context first needs the kernel (csrc/sab_module.cu).  This returns a copy of the
 * image of <method, packed, width> specialised for `gap_constant` -- what `cuobjdump -sass` of
 * the running kernel would show -- and the number of instruction words that carry the constant
 * (< 0: error).  out == NULL only reports *size and the kernel's (mangled) name.  Host only. */
int sab_kernel_image(int method, int packed, int width, uint32_t gap_constant, unsigned char *out,
		     size_t cap, size_t *size, char *kernel_name, size_t name_cap);

/* Host-only self-test of the upload packing: packs the set as sab_upload would (length binning, the
 * 2-/5-bit subject bit stream, the lane-interleaved query words) and reads every letter back with the
 * kernels' addressing.  Returns the number of letters that differ (0 = good), < 0 on errors;
 * *bits_per_letter = subject packing chosen (2: the set uses at most four letters, else 5). */
int sab_selftest_packing(const sab_config *cfg, const uint8_t *seqs, const sab_meta *meta, int32_t num,
			 int32_t *bits_per_letter);

/*
 * Every GPU of the box from one host process (what cuda_align needs: it is called once, on
 * the main thread, src/main.c:31): one context and one host thread per device, each device
 * holds a replica of the sequence set and pulls row strips (costliest first) from a cursor the
 * devices share, writing them into the caller's matrix.  No collective.
 * ndev > 0: exactly that many devices.  ndev <= 0: as many as the job can keep busy -- contexts
 * are created lazily, in parallel, when sab_multi_upload / sab_multi_filter know the job's
 * size (a context costs the host a few hundred ms; a small job is faster on one device).
 * Stats: counts are summed over the devices, times are the slowest device's.  The progress
 * callback is serialised and reports pairs over the whole job.
 */
typedef struct sab_multi sab_multi;
int sab_multi_create(sab_multi **m, const sab_config *cfg, int32_t ndev);
void sab_multi_destroy(sab_multi *m);
const char *sab_multi_last_error(const sab_multi *m);
int sab_multi_device_count(const sab_multi *m);
int sab_multi_upload(sab_multi *m, const uint8_t *seqs, const sab_meta *meta, int32_t num,
		     int32_t max);
int sab_multi_align(sab_multi *m, int32_t *matrix, int triangular, sab_progress_fn progress,
		    void *user);
/* sab_filter on the devices of `m`: phase A sharded over the devices, phase B on the host. */
int sab_multi_filter(sab_multi *m, const uint8_t *seqs, const sab_meta *meta, int32_t num,
		     float threshold, uint8_t *lost);
int sab_multi_get_stats(const sab_multi *m, sab_stats *out);

/*
 * Redundancy filter decisions (src/bio/filter.c:14-62, the deterministic
 * sequential semantics): lost[j] = 1 iff some KEPT i<j has
 * (float)matches/(float)min(len_i,len_j) >= threshold over the common prefix.
 * Works on the raw letters given here (independent of sab_upload).  Returns the
 * number of lost sequences (>= 0) or a negative sab_status.
 */
int sab_filter(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
	       float threshold, uint8_t *lost);
/*
 * The two phases of sab_filter, for hosts that spread the filter over several devices or
 * processes.  Phase A (device): every pair i<j of shard `shard` of `nshards` (tiles of 32 j's,
 * dealt out so the shards' loads stay level) with (float)matches/(float)min(len) >= threshold,
 * written as (j << 32 | i) in no particular order; *count receives how many.  More than `cap`
 * -> SAB_ERR_OVERFLOW (use sab_filter, whose panel-sequential path handles any density).
 * Phase B (host only): the order-dependent recurrence of filter.c:35-55 over the concatenated
 * lists of all shards; returns the number of lost sequences.
 */
int sab_filter_matches(sab_ctx *ctx, const uint8_t *seqs, const sab_meta *meta, int32_t num,
		       float threshold, int32_t shard, int32_t nshards, uint64_t *pairs, int64_t cap,
		       int64_t *count);
int sab_filter_resolve(int32_t num, const uint64_t *pairs, int64_t count, uint8_t *lost);

/* Built-in substitution tables (host only, no device needed). */
int sab_matrix_count(void);
const char *sab_matrix_name(int index);
/* Fills SEQ_LUT / SUB_MAT for table `name` (case-insensitive, src/bio/matrices.c:47).
 * Returns SAB_OK or SAB_ERR_ARG. */
int sab_matrix_lookup(const char *name, int32_t seq_lut[128], int32_t sub_mat[24 * 24]);

/*
 * Output format after the path (host only, no device needed): writes the HDF5 file of
 * flush_hdf5() (src/io/format/hdf5.c:14-202) -- /sequences = dim variable-length C strings,
 * /similarity_matrix = [dim,dim] int32 LE, contiguous when dim <= 256, else chunked
 * (sab_h5_chunk_dim) with deflate level `compression` (0..9, 0 = none) -- without libhdf5.
 * `matrix` is in either layout sab_align fills; a triangular one is expanded to the full
 * symmetric matrix (diagonal 0) chunk by chunk.  `progress` gets (chunk rows done, total).
 */
typedef void (*sab_h5_progress_fn)(int64_t done, int64_t total, void *user);
int sab_h5_write(const char *path, const char *const *seqs, size_t dim, const int32_t *matrix,
		 int triangular, unsigned compression, sab_h5_progress_fn progress, void *user);
/* Same with an explicit chunk edge (0 = contiguous, < 0 = the reference's rule). */
int sab_h5_write_chunked(const char *path, const char *const *seqs, size_t dim, const int32_t *matrix,
			 int triangular, unsigned compression, int chunk_dim,
			 sab_h5_progress_fn progress, void *user);
/* Chunk edge the reference's rule evaluates to (hdf5.c:68-84); 0 = contiguous. */
int sab_h5_chunk_dim(size_t dim);

/*
 * Input formats before the path (host only, no device needed): the text parsers that build the
 * sequence store sab_upload takes.  Same accept/reject decisions, same store and the same
 * messages as the reference's parsers; one pass, table driven, in place.
 *
 * `file` holds the whole input file (`bytes` of it) and is rewritten IN PLACE, front to back,
 * into the store layout of struct input (src/io/input.h:6-11): upper-cased letters, blanks,
 * line ends (and, DSV, quotes) dropped, one NUL after every sequence.  The store never outgrows
 * the text it came from, with one exception the reference shares: a headerless single-column
 * DSV file whose last row has no line end needs file[bytes] for the final NUL.
 * seq_lut = SEQ_LUT of the selected matrix (every stored letter has seq_lut[c] >= 0); gap_pen =
 * GAP_PEN (<= 0), which bounds the sequence length (sequence_length_limit, src/io/input.c:15-19).
 * On SAB_ERR_PARSE `err` (capacity err_cap, may be NULL) receives the message the reference
 * prints for that input and the content of `file` is unspecified.
 */
enum sab_source { SAB_SRC_NONE = 0, SAB_SRC_FASTA = 1, SAB_SRC_DSV = 2 };
/* Format of a file by its extension (case-insensitive, without the dot): fasta fa fas fna ffn faa
 * frn mpfa (src/io/source/fasta.c:12-14); csv ',' tsv '\t' ssv ';' psv '|' (src/io/source/dsv.c:14-19,
 * *delimiter receives the field separator).  SAB_SRC_NONE: no parser takes it. */
int sab_source_format(const char *ext, uint8_t *delimiter);
/* parse_fasta, src/io/source/fasta.c:26-85: a record = header line ('>' up to the first line end)
 * + letters up to the next '>'.  *num = records, *max = longest sequence. */
int sab_parse_fasta(uint8_t *file, size_t bytes, const int32_t seq_lut[128], int32_t gap_pen,
		    int32_t *num, int32_t *max, char *err, size_t err_cap);
/* First row of a DSV file (src/io/source/dsv.c:94-144): *columns = fields of the row, *seq_column =
 * the first one named sequence / seq / protein / dna / rna / amino / peptide / chain (any case), or
 * -1 -- the reference then ASKS which column holds the sequences and whether the first row is a
 * header at all (dsv.c:146-160); the host does the asking and passes the answer to sab_parse_dsv.
 * `names` (may be NULL) receives the field texts, each NUL-terminated, back to back.
 * SAB_ERR_PARSE "First row has empty column" as dsv.c:106-112. */
int sab_dsv_header(const uint8_t *file, size_t bytes, uint8_t delimiter, int32_t *columns,
		   int32_t *seq_column, char *names, size_t names_cap, char *err, size_t err_cap);
/* The row loop of parse_dsv (src/io/source/dsv.c:163-227): field `seq_column` of every row is a
 * sequence; quoting as dsv_field (dsv.c:26-57: a delimiter or line end inside "..." belongs to the
 * field, "" is a quote); rows with fewer or more than the first row's columns are errors.
 * first_row_is_data != 0: there is no header row (the reference's "No header line" answer). */
int sab_parse_dsv(uint8_t *file, size_t bytes, uint8_t delimiter, int32_t seq_column,
		  int first_row_is_data, const int32_t seq_lut[128], int32_t gap_pen, int32_t *num,
		  int32_t *max, char *err, size_t err_cap);
/* meta[k] = (offset, length) of the k-th NUL-terminated sequence of a store written by
 * sab_parse_* (src/io/input.c:70-81).  Returns the store's size in bytes (terminators included). */
int64_t sab_build_meta(const uint8_t *seqs, int32_t num, sab_meta *meta);

/*
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Smith-Waterman scores (src/bio/method/sw.c:15-61 arithmetic, int32) of a caller-given list of pairs,
 * computed WITHOUT any allocation: no cudaMalloc / cudaFree / cudaHostAlloc, no allocating library call, no
 * host <-> device copy.  Every buffer is a DEVICE buffer of the current device, owned by the caller:
 *   d_seqs, d_meta   device copies of struct input's `seqs` / `meta` (src/io/input.h:6-11), `num` sequences,
 *                    none longer than `max_len`
 *   d_pair_i/j       `count` pairs of sequence indices (i != j, either order)
 *   d_scores         out: d_scores[k] = score(d_pair_i[k], d_pair_j[k]) -- what sab_align writes to
 *                    matrix[j(j-1)/2 + i]; INT32_MIN for a bad index, a length outside 1..max_len or a letter
 *                    outside the matrix alphabet
 * cfg: method must be SAB_SW; gap_open, gap_ext, seq_lut, sub_mat as for sab_create (device, panel, flags
 * unused).  The kernel is enqueued on `stream` (a cudaStream_t; NULL = the default stream) and the call
 * returns without waiting for it, like a kernel launch.
 *
 * The kernel's scratch (two DP rows per thread) is ONE file-scope device array of SAB_NOALLOC_SCRATCH_BYTES
 * that every GPU which loads the library holds until exit.  Limit: max_len <= SAB_NOALLOC_MAX_LEN, beyond it
 * SAB_ERR_ARG and a message in `err` (use sab_align).  All calls on a device share that array: calls that
 * OVERLAP on one GPU (two streams, concurrent replays of a captured graph) race on it and must be serialised
 * by the caller; calls on one stream are ordered and safe.  The array keeps the previous call's contents;
 * nothing in it needs to start at zero (every thread writes a pair's border row before reading it).
 * The first call on a device looks the array's address up (cudaGetSymbolAddress: may load the library's
 * device code); a call with count == 0 does only that -- make it once per device before capturing a graph.
 * One thread per pair: meant for lists of pairs, not for all-vs-all throughput.
 */
#define SAB_NOALLOC_SCRATCH_BYTES (16u << 20)
#define SAB_NOALLOC_MAX_LEN 4095
int sab_sw_pairs_noalloc(const sab_config *cfg, const uint8_t *d_seqs, const sab_meta *d_meta, int32_t num,
			 int32_t max_len, const int32_t *d_pair_i, const int32_t *d_pair_j, int64_t count,
			 int32_t *d_scores, void *stream, char *err, size_t err_cap);

#ifdef __cplusplus
}
#endif

#endif /* SEQALIGN_B200_H */
