// sab_tables.cpp -- built-in substitution tables (host only, no device needed).
//
// Same data the reference selects with `-m` (reference: src/bio/matrices.c:44-57
// parse_matrix; tables src/generated/matrices.h:4-9), stored here as int8 upper
// triangles (submat_tables.inc, written by tools/extract_matrices.py) and expanded
// to the SEQ_LUT[128] / SUB_MAT[24][24] int32 form the boundary uses.
#include "../../include/seqalign_b200.h"

#include <strings.h>

#include "submat_tables.inc"

extern "C" {

int sab_matrix_count(void)
{
	return SAB_N_AMINO + SAB_N_NUCLEO;
}

const char *sab_matrix_name(int index)
{
	if (index < 0 || index >= sab_matrix_count())
		return nullptr;
	return SAB_TABLE_INDEX[index].name;
}

int sab_matrix_lookup(const char *name, int32_t seq_lut[128], int32_t sub_mat[24 * 24])
{
	if (!name || !seq_lut || !sub_mat)
		return SAB_ERR_ARG;
	for (int k = 0; k < sab_matrix_count(); k++) {
		if (strcasecmp(name, SAB_TABLE_INDEX[k].name) != 0) // matrices.c:47
			continue;
		const int dim = SAB_TABLE_INDEX[k].dim;
		const char *alphabet = k < SAB_N_AMINO ? SAB_ALPHABET_AMINO : SAB_ALPHABET_NUCLEO;
		for (int c = 0; c < 128; c++)
			seq_lut[c] = -1;
		for (int a = 0; a < dim; a++)
			seq_lut[(unsigned char)alphabet[a]] = a;
		for (int i = 0; i < 24 * 24; i++)
			sub_mat[i] = 0;
		const signed char *p = SAB_TABLE_BLOB + SAB_TABLE_INDEX[k].first;
		for (int a = 0; a < dim; a++)
			for (int b = a; b < dim; b++) {
				const int v = *p++;
				sub_mat[a * 24 + b] = v;
				sub_mat[b * 24 + a] = v;
			}
		return SAB_OK;
	}
	return SAB_ERR_ARG;
}

} // extern "C"
