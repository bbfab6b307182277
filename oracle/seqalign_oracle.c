/*
 * oracle/seqalign_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's score-only all-vs-all hot path
 * (jakovdev/SequenceAligner).  It is the checker for the CUDA product path and
 * the "port" CPU baseline of bench.py.  Nothing under sequencealigner_b200/ may
 * import, link or execute it; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity pin: the reference ships no golden vectors, so
 * this restatement is pinned against the UNMODIFIED reference CPU sources
 * compiled into oracle/_ref/seqalign (see oracle/Makefile) -- in this container
 * by tests/test_oracle.py, and everywhere through the committed
 * fixtures under tests/golden/ that were produced by that binary
 * (tools/make_golden.py).
 *
 * The recurrences are restated with two rolling rows instead of the
 * reference's full (len+1)^2 tables; the arithmetic (int32, SCORE_MIN sentinel,
 * operand order of the substitution lookup) follows the reference exactly:
 *
 *   oracle_nw  <- src/bio/method/nw.c:5-42   (Needleman-Wunsch, linear gap)
 *   oracle_ga  <- src/bio/method/ga.c:10-68  (Gotoh, global affine)
 *   oracle_sw  <- src/bio/method/sw.c:5-62   (Smith-Waterman, local affine)
 *   oracle_align_all <- src/bio/align.c:21-72 + src/io/output.c:68-84
 *   oracle_filter    <- src/bio/filter.c:14-89 (sequential, i.e. -T 1, semantics)
 *
 * Naming follows the reference: sequence j is the "outer" one (mapped through
 * SEQ_LUT into ind[], length len1), sequence i<j the "inner" one (raw letters,
 * length len2).  Gap values are stored <= 0 (align.c:127).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SUB_DIM 24
#define SCORE_MIN (INT32_MIN / 2) /* bio/align.h:19 */

typedef struct {
	int32_t off;
	int32_t len;
} oracle_meta; /* bio/align.h:6-9 */

typedef struct {
	int32_t method; /* 0 nw, 1 ga, 2 sw */
	int32_t gap_pen, gap_open, gap_ext; /* <= 0 */
	int32_t seq_lut[128];
	int32_t sub_mat[SUB_DIM * SUB_DIM];
} oracle_config;

static inline int32_t max2(int32_t a, int32_t b)
{
	return a > b ? a : b;
}

/* nw.c:5-42.  H[r][c] over rows = inner seq (len2), cols = outer seq (len1).
 * Substitution operand order: SUB_MAT[inner letter][outer letter] (nw.c:22-27). */
static int32_t oracle_nw(const oracle_config *cfg, int32_t len1, const int32_t *ind,
			 int32_t len2, const uint8_t *seq2, int32_t *work)
{
	const int32_t g = cfg->gap_pen;
	int32_t *h = work; /* len1 + 1 */
	for (int32_t c = 0; c <= len1; c++)
		h[c] = c * g;
	for (int32_t r = 1; r <= len2; r++) {
		const int32_t *sub = cfg->sub_mat + SUB_DIM * cfg->seq_lut[seq2[r - 1]];
		int32_t diag = h[0];
		int32_t left = r * g;
		h[0] = left;
		for (int32_t c = 1; c <= len1; c++) {
			int32_t up = h[c];
			int32_t v = diag + sub[ind[c - 1]];
			v = max2(up + g, v);
			v = max2(left + g, v);
			diag = up;
			h[c] = v;
			left = v;
		}
	}
	return h[len1];
}

/* ga.c:10-68 and sw.c:5-62 share the three-state cell; `local` selects the
 * zero floor, zero borders and running maximum of sw.c.
 * Substitution operand order: SUB_MAT[outer letter][inner letter] (ga.c:46, sw.c:39). */
static int32_t oracle_affine(const oracle_config *cfg, int local, int32_t len1,
			     const int32_t *ind, int32_t len2, const uint8_t *seq2,
			     int32_t *work)
{
	const int32_t o = cfg->gap_open, e = cfg->gap_ext;
	int32_t *m = work; /* match row        */
	int32_t *y = work + (len1 + 1); /* gap_y row (vertical) */
	m[0] = 0;
	y[0] = SCORE_MIN;
	{ /* row 0: ga.c:26-30 / sw.c:21-24 */
		int32_t x = SCORE_MIN;
		for (int32_t c = 1; c <= len1; c++) {
			if (local) {
				m[c] = 0;
			} else {
				x = max2(m[c - 1] + o, x + e);
				m[c] = x;
			}
			y[c] = SCORE_MIN;
		}
	}
	int32_t best = 0;
	for (int32_t r = 1; r <= len2; r++) {
		const int32_t c2 = cfg->seq_lut[seq2[r - 1]];
		int32_t diag = m[0];
		/* column 0: ga.c:32-38 / sw.c:26-30 */
		if (local) {
			m[0] = 0;
			y[0] = SCORE_MIN;
		} else {
			y[0] = max2(m[0] + o, y[0] + e);
			m[0] = y[0];
		}
		int32_t x = SCORE_MIN; /* gap_x[row][0] */
		int32_t left = m[0];
		for (int32_t c = 1; c <= len1; c++) {
			int32_t s = cfg->sub_mat[SUB_DIM * ind[c - 1] + c2];
			int32_t d = diag + s;
			x = max2(left + o, x + e);
			int32_t yy = max2(m[c] + o, y[c] + e);
			int32_t v = local ? max2(d, 0) : d;
			v = max2(x, v);
			v = max2(yy, v);
			diag = m[c];
			m[c] = v;
			y[c] = yy;
			left = v;
			if (local)
				best = max2(best, v);
		}
	}
	return local ? best : m[len1];
}

/* Scores one (i<j) pair exactly as `ALIGN->method(l1, l2, s2, ind, table)` does
 * in align.c:55.  seq_j / seq_i are raw uppercase ASCII letters. */
int32_t oracle_pair(const oracle_config *cfg, const uint8_t *seq_j, int32_t len_j,
		    const uint8_t *seq_i, int32_t len_i)
{
	int32_t *ind = malloc(sizeof(int32_t) * (size_t)(len_j > 0 ? len_j : 1));
	int32_t *work = malloc(sizeof(int32_t) * 2 * ((size_t)len_j + 1));
	for (int32_t k = 0; k < len_j; k++)
		ind[k] = cfg->seq_lut[seq_j[k]];
	int32_t s;
	if (cfg->method == 0)
		s = oracle_nw(cfg, len_j, ind, len_i, seq_i, work);
	else
		s = oracle_affine(cfg, cfg->method == 2, len_j, ind, len_i, seq_i, work);
	free(work);
	free(ind);
	return s;
}

/* align.c:21-72 with output_fill (output.c:68-84) folded in.
 * matrix: full  -> matrix[num*i+j] = matrix[num*j+i] = score, diagonal untouched;
 *         packed-triangular -> matrix[j(j-1)/2 + i] = score, 0 <= i < j < num.
 * Only columns j with first_j <= j < last_j are computed (sampling for the
 * timed CPU baseline); pass 1, num for everything. Returns number of cells. */
int64_t oracle_align_all(const oracle_config *cfg, const uint8_t *seqs,
			 const oracle_meta *meta, int32_t num, int32_t max_len,
			 int32_t *matrix, int32_t triangular, int32_t first_j,
			 int32_t last_j)
{
	int64_t cells = 0;
	if (first_j < 1)
		first_j = 1;
	if (last_j > num)
		last_j = num;
#pragma omp parallel reduction(+ : cells)
	{
		int32_t *ind = malloc(sizeof(int32_t) * (size_t)(max_len > 0 ? max_len : 1));
		int32_t *work = malloc(sizeof(int32_t) * 2 * ((size_t)max_len + 1));
#pragma omp for schedule(dynamic)
		for (int32_t j = first_j; j < last_j; j++) {
			const uint8_t *s1 = seqs + meta[j].off;
			const int32_t l1 = meta[j].len;
			for (int32_t k = 0; k < l1; k++)
				ind[k] = cfg->seq_lut[s1[k]];
			for (int32_t i = 0; i < j; i++) {
				const uint8_t *s2 = seqs + meta[i].off;
				const int32_t l2 = meta[i].len;
				int32_t s;
				if (cfg->method == 0)
					s = oracle_nw(cfg, l1, ind, l2, s2, work);
				else
					s = oracle_affine(cfg, cfg->method == 2, l1, ind, l2,
							  s2, work);
				cells += (int64_t)l1 * l2;
				if (!matrix)
					continue;
				if (triangular) {
					matrix[(int64_t)j * (j - 1) / 2 + i] = s;
				} else {
					matrix[(int64_t)num * i + j] = s;
					matrix[(int64_t)num * j + i] = s;
				}
			}
		}
		free(work);
		free(ind);
	}
	return cells;
}

/* filter.c:14-62, sequential (-T 1) semantics: sequence j is lost iff some KEPT
 * i<j has prefix identity matches/min(len) >= threshold (float32 divide+compare).
 * lost[] must hold num bytes; returns the number of lost sequences. */
int32_t oracle_filter(const uint8_t *seqs, const oracle_meta *meta, int32_t num,
		      float threshold, uint8_t *lost)
{
	memset(lost, 0, (size_t)num);
	if (threshold <= 0.0f)
		return 0;
	int32_t nlost = 0;
	for (int32_t j = 1; j < num; j++) {
		const uint8_t *s1 = seqs + meta[j].off;
		for (int32_t i = 0; i < j; i++) {
			if (lost[i])
				continue;
			const uint8_t *s2 = seqs + meta[i].off;
			int32_t ml = meta[j].len < meta[i].len ? meta[j].len : meta[i].len;
			int32_t matches = 0;
			for (int32_t k = 0; k < ml; k++)
				matches += s1[k] == s2[k];
			if ((float)matches / (float)ml >= threshold) {
				lost[j] = 1;
				nlost++;
				break;
			}
		}
	}
	return nlost;
}

/* A list of (i<j) pairs scored as align.c:55 scores them (sampled parity at sizes the whole
 * matrix cannot be recomputed at): out[k] = score(is[k], js[k]).  OpenMP over the list. */
void oracle_pairs(const oracle_config *cfg, const uint8_t *seqs, const oracle_meta *meta,
		  int64_t n_pairs, const int32_t *is, const int32_t *js, int32_t *out)
{
#pragma omp parallel for schedule(dynamic, 64)
	for (int64_t k = 0; k < n_pairs; k++)
		out[k] = oracle_pair(cfg, seqs + meta[js[k]].off, meta[js[k]].len, seqs + meta[is[k]].off,
				     meta[is[k]].len);
}
