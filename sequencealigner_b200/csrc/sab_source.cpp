// sab_source.cpp -- the text formats in front of the path (host only): FASTA and delimiter-separated
// files -> the sequence store of struct input (src/io/input.h:6-11).
//
// Behaviour = the reference's parsers (src/io/source/fasta.c:16-85, src/io/source/dsv.c:26-230, the
// meta pass of src/io/input.c:70-81): same accepted inputs, same store, same message for every
// rejected input (tests/test_sources.py replays both against the reference binary).  The code is
// organised differently: every input byte is classified ONCE through a 256-entry table built from
// the matrix's SEQ_LUT (letter to store, byte to drop, record end, corrupt, not in the alphabet),
// so the per-byte work of the copy loop is one table load and one predictable branch instead of
// toupper() + five compares + the LUT load; FASTA lines that hold nothing but letters are found by an
// OR-scan of the class flags and block-copied, and header lines are skipped with memchr.
#include "../../include/seqalign_b200.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <strings.h>

namespace {

constexpr int32_t kS32Max = 2147483647;
constexpr int32_t kSeqLenMax = (kS32Max - 1) / 2; // SEQ_LEN_MAX, src/bio/align.h:21-23

int fail(char *err, size_t cap, const char *fmt, ...)
{
	if (err && cap) {
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(err, cap, fmt, ap);
		va_end(ap);
	}
	return SAB_ERR_PARSE;
}

// What the copy loop does with a byte.
enum : uint8_t {
	B_LETTER = 0, // store its upper-case form
	B_DROP = 1, // blank / line end (DSV: also a quote)
	B_END = 2, // FASTA: '>' ends the record
	B_CORRUPT = 3, // NUL or >= 0x80                    "Sequence #n is corrupted"
	B_FOREIGN = 4, // not a letter of the matrix        "Sequence #n is invalid"
};

struct ByteClass {
	uint8_t kind[256];
	uint8_t upper[256];
	uint8_t flags[256]; // 0 = upper-case letter, kLower = lower-case letter, kOther = anything else
	static constexpr uint8_t kLower = 1, kOther = 2;

	// fasta.c:52-64 / dsv.c:186-199: upper-case first, then drop, then corrupt, then the alphabet.
	ByteClass(const int32_t lut[128], const bool dsv)
	{
		for (int c = 0; c < 256; c++) {
			const int u = (c >= 'a' && c <= 'z') ? c - 32 : c;
			upper[c] = (uint8_t)u;
			uint8_t k;
			if (u == '\r' || u == '\n' || u == ' ' || (dsv && u == '"'))
				k = B_DROP;
			else if (u == 0 || u > 127)
				k = B_CORRUPT;
			else
				k = lut[u] < 0 ? B_FOREIGN : B_LETTER;
			kind[c] = k;
		}
		if (!dsv)
			kind['>'] = B_END; // tested before anything else (fasta.c:51)
		for (int c = 0; c < 256; c++)
			flags[c] = kind[c] != B_LETTER ? kOther : upper[c] != c ? kLower : 0;
	}

	// OR of the flags of [p, p + n): tells whether the run can be block-copied
	uint8_t scan(const uint8_t *p, const size_t n) const
	{
		uint8_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
		size_t i = 0;
		for (; i + 4 <= n; i += 4) {
			a0 |= flags[p[i]];
			a1 |= flags[p[i + 1]];
			a2 |= flags[p[i + 2]];
			a3 |= flags[p[i + 3]];
		}
		for (; i < n; i++)
			a0 |= flags[p[i]];
		return a0 | a1 | a2 | a3;
	}
};

inline bool line_end(const uint8_t c)
{
	return c == '\n' || c == '\r';
}

// sequence_length_limit, src/io/input.c:15-19
inline bool length_ok(const int32_t len, const int32_t gap_pen)
{
	const int64_t gap = -(int64_t)gap_pen;
	return gap ? len <= kSeqLenMax / gap : len <= kSeqLenMax;
}

// Running totals of a store being written; the checks every finished sequence goes through
// (fasta.c:64-79 == dsv.c:200-215).
struct StoreWriter {
	uint8_t *w;
	int32_t num = 0, longest = 0;
	int64_t total = 0;

	// `terminate` = false leaves the NUL to the caller (DSV: the byte may still be unread input)
	int close_sequence(const int32_t len, const int32_t gap_pen, char *err, const size_t cap,
			   const bool terminate = true)
	{
		if (!len)
			return fail(err, cap, "Sequence #%d is empty", num);
		if (!length_ok(len, gap_pen))
			return fail(err, cap, "Sequence #%d exceeds length limits", num);
		if (total + len + 1 > kS32Max)
			return fail(err, cap, "Length overflow after %d sequences", num);
		if (len > longest)
			longest = len;
		total += len + 1;
		if (terminate)
			*w++ = 0;
		return SAB_OK;
	}
};

// One DSV field starting at *cur (dsv.c:26-57): runs to the next delimiter or line end outside
// quotes, "" inside quotes is a literal quote; an enclosing pair of quotes is not part of the text;
// the delimiter after the field is consumed, a line end is not.
const uint8_t *take_field(const uint8_t **cur, const uint8_t *end, const uint8_t delim, int32_t *flen)
{
	const uint8_t *p = *cur;
	const uint8_t *first = p;
	bool in_quotes = false;
	for (; p < end; p++) {
		const uint8_t c = *p;
		if (c == '"') {
			if (in_quotes && p + 1 < end && p[1] == '"')
				p++; // "" -> skip both
			else
				in_quotes = !in_quotes;
		} else if (!in_quotes && (c == delim || line_end(c))) {
			break;
		}
	}
	int32_t len = (int32_t)(p - first);
	if (len >= 2 && first[0] == '"' && first[len - 1] == '"') {
		first++;
		len -= 2;
	}
	*flen = len;
	*cur = (p < end && *p == delim) ? p + 1 : p;
	return first;
}

// Fields of the first row (dsv.c:59-78).
int32_t count_columns(const uint8_t *p, const uint8_t *end, const uint8_t delim)
{
	int32_t n = 1;
	bool in_quotes = false;
	while (p < end) {
		const uint8_t c = *p;
		if (c == '"') {
			if (in_quotes && p + 1 < end && p[1] == '"') {
				p += 2;
				continue;
			}
			in_quotes = !in_quotes;
		} else if (c == delim && !in_quotes) {
			n++;
		}
		if (!in_quotes && line_end(c))
			break;
		p++;
	}
	return n;
}

const char *const kFastaExt[] = { "fasta", "fa", "fas", "fna", "ffn", "faa", "frn", "mpfa" };
const struct {
	const char *ext;
	uint8_t delim;
} kDsvExt[] = { { "csv", ',' }, { "tsv", '\t' }, { "ssv", ';' }, { "psv", '|' } };
const char *const kSeqColumnNames[] = { "sequence", "seq", "protein", "dna", "rna", "amino", "peptide", "chain" };

} // namespace

extern "C" {

int sab_source_format(const char *ext, uint8_t *delimiter)
{
	if (!ext)
		return SAB_SRC_NONE;
	for (const char *e : kFastaExt)
		if (strcasecmp(e, ext) == 0)
			return SAB_SRC_FASTA;
	for (const auto &d : kDsvExt)
		if (strcasecmp(d.ext, ext) == 0) {
			if (delimiter)
				*delimiter = d.delim;
			return SAB_SRC_DSV;
		}
	return SAB_SRC_NONE;
}

int sab_parse_fasta(uint8_t *file, size_t bytes, const int32_t seq_lut[128], int32_t gap_pen, int32_t *num,
		    int32_t *max, char *err, size_t err_cap)
{
	if (!file || !seq_lut || !num || !max || gap_pen > 0 || bytes > (size_t)kS32Max)
		return SAB_ERR_ARG;
	const ByteClass cls(seq_lut, false);
	const uint8_t *p = file, *const end = file + bytes;
	if (p >= end || *p != '>')
		return fail(err, err_cap, "Data before first header");
	StoreWriter st{ file };
	while (p < end) {
		// header line: up to the first line end, then every line end that follows (fasta.c:39-46)
		const uint8_t *nl = (const uint8_t *)memchr(p, '\n', (size_t)(end - p));
		const uint8_t *cr = (const uint8_t *)memchr(p, '\r', (size_t)((nl ? nl : end) - p));
		p = cr ? cr : nl ? nl : end;
		while (p < end && line_end(*p))
			p++;
		if (p >= end)
			return fail(err, err_cap, "Last header has no data");
		st.num++;
		uint8_t *const w0 = st.w;
		uint8_t *w = w0;
		// The record's text, line by line.  A line made of letters only -- nearly every line of a real
		// file -- is block-copied (memmove when already upper case); any other line goes byte by byte,
		// which is where blanks are dropped, '>' ends the record and the first bad byte is reported.
		for (bool open = true; open && p < end;) {
			const uint8_t *nl = (const uint8_t *)memchr(p, '\n', (size_t)(end - p));
			const uint8_t *const stop = nl ? nl : end;
			size_t n = (size_t)(stop - p);
			if (n && stop[-1] == '\r')
				n--; // the CR of a CRLF line end is dropped like the LF
			const uint8_t mix = cls.scan(p, n);
			if (mix == 0) {
				memmove(w, p, n); // w <= p: the store never overtakes the text
				w += n;
			} else if (mix == ByteClass::kLower) {
				for (size_t i = 0; i < n; i++)
					w[i] = cls.upper[p[i]];
				w += n;
			} else {
				for (; p < stop; p++) {
					const uint8_t k = cls.kind[*p];
					if (k == B_LETTER) {
						*w++ = cls.upper[*p];
					} else if (k == B_DROP) {
						continue;
					} else if (k == B_END) {
						open = false;
						break;
					} else {
						return fail(err, err_cap,
							    k == B_CORRUPT ? "Sequence #%d is corrupted" :
									     "Sequence #%d is invalid",
							    st.num);
					}
				}
			}
			if (open)
				p = stop < end ? stop + 1 : end;
		}
		st.w = w;
		const int rc = st.close_sequence((int32_t)(w - w0), gap_pen, err, err_cap);
		if (rc != SAB_OK)
			return rc;
	}
	*num = st.num;
	*max = st.longest;
	return SAB_OK;
}

int sab_dsv_header(const uint8_t *file, size_t bytes, uint8_t delimiter, int32_t *columns, int32_t *seq_column,
		   char *names, size_t names_cap, char *err, size_t err_cap)
{
	if (!file || !columns || !seq_column || bytes > (size_t)kS32Max)
		return SAB_ERR_ARG;
	const uint8_t *p = file, *const end = file + bytes;
	const int32_t cols = count_columns(p, end, delimiter);
	int32_t found = -1;
	size_t used = 0;
	for (int32_t col = 0; col < cols; col++) {
		int32_t flen;
		const uint8_t *f = take_field(&p, end, delimiter, &flen);
		if (!flen)
			return fail(err, err_cap, "First row has empty column");
		// strcasecmp on the NUL-terminated copy (dsv.c:130-139): a field with an embedded NUL
		// compares as its part before the NUL
		const size_t vis = strnlen((const char *)f, (size_t)flen);
		if (found < 0)
			for (const char *key : kSeqColumnNames)
				if (strlen(key) == vis && strncasecmp(key, (const char *)f, vis) == 0) {
					found = col;
					break;
				}
		if (names) {
			if (used + vis + 1 > names_cap)
				return SAB_ERR_ARG;
			memcpy(names + used, f, vis);
			names[used + vis] = 0;
			used += vis + 1;
		}
	}
	*columns = cols;
	*seq_column = found;
	return SAB_OK;
}

int sab_parse_dsv(uint8_t *file, size_t bytes, uint8_t delimiter, int32_t seq_column, int first_row_is_data,
		  const int32_t seq_lut[128], int32_t gap_pen, int32_t *num, int32_t *max, char *err,
		  size_t err_cap)
{
	if (!file || !seq_lut || !num || !max || gap_pen > 0 || bytes > (size_t)kS32Max)
		return SAB_ERR_ARG;
	const ByteClass cls(seq_lut, true);
	const uint8_t *p = file, *const end = file + bytes;
	const int32_t cols = count_columns(p, end, delimiter);
	if (seq_column < 0 || seq_column >= cols)
		return SAB_ERR_ARG;
	int32_t flen;
	// The first row is always read as a row of names first (dsv.c:103-126), so an empty field in it
	// is an error even when the row then turns out to be data (dsv.c:151-152 rewinds to it).
	for (int32_t col = 0; col < cols; col++) {
		take_field(&p, end, delimiter, &flen);
		if (!flen)
			return fail(err, err_cap, "First row has empty column");
	}
	if (first_row_is_data)
		p = file;
	StoreWriter st{ file };
	for (;;) {
		while (p < end && line_end(*p))
			p++;
		if (p >= end)
			break;
		st.num++;
		for (int32_t col = 0; col < seq_column; col++) {
			take_field(&p, end, delimiter, &flen);
			if (p >= end || line_end(*p))
				return fail(err, err_cap, "DSV row #%d has no sequence column", st.num);
		}
		const uint8_t *f = take_field(&p, end, delimiter, &flen);
		if (!flen)
			return fail(err, err_cap, "Sequence #%d is empty", st.num);
		uint8_t *const w0 = st.w;
		uint8_t *w = w0;
		for (int32_t i = 0; i < flen; i++) {
			const uint8_t k = cls.kind[f[i]];
			if (k == B_LETTER)
				*w++ = cls.upper[f[i]];
			else if (k != B_DROP)
				return fail(err, err_cap,
					    k == B_CORRUPT ? "Sequence #%d is corrupted" : "Sequence #%d is invalid",
					    st.num);
		}
		st.w = w;
		// In a headerless one-column file the store catches up with the text: the terminator of this
		// sequence would land on the line end that closes its row, so it is written once the row's
		// remaining checks are done and its line end has been passed.  (The reference writes it first, dsv.c:215, and cannot notice: it never
		// gets as far as parsing such a file, see tests/test_sources.py.)
		const int rc = st.close_sequence((int32_t)(w - w0), gap_pen, err, err_cap, false);
		if (rc != SAB_OK)
			return rc;
		for (int32_t col = seq_column + 1; col < cols; col++) {
			if (p >= end || line_end(*p))
				return fail(err, err_cap, "DSV row #%d has too few columns", st.num);
			take_field(&p, end, delimiter, &flen);
		}
		if (p < end && !line_end(*p))
			return fail(err, err_cap, "DSV row #%d has too many columns", st.num);
		while (p < end && line_end(*p))
			p++;
		*st.w++ = 0;
	}
	*num = st.num;
	*max = st.longest;
	return SAB_OK;
}

int64_t sab_build_meta(const uint8_t *seqs, int32_t num, sab_meta *meta)
{
	if (!seqs || !meta || num < 0)
		return SAB_ERR_ARG;
	const uint8_t *p = seqs;
	for (int32_t k = 0; k < num; k++) {
		const size_t len = strlen((const char *)p);
		meta[k].off = (int32_t)(p - seqs);
		meta[k].len = (int32_t)len;
		p += len + 1;
	}
	return (int64_t)(p - seqs);
}

} // extern "C"
