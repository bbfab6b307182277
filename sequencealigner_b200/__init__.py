"""sequencealigner_b200 -- B200-native (sm_100a) score-only all-vs-all pairwise
alignment (Needleman-Wunsch, Gotoh, Smith-Waterman), drop-in for that one path of
jakovdev/SequenceAligner.  The product is libseqalign_b200.so (C ABI:
include/seqalign_b200.h); this package is the Python host binding around it."""
from .aligner import (AlignConfig, Aligner, MultiAligner, align_all, matrix_lookup,  # noqa: F401
                      matrix_names)
from .seqstore import ParseError, SeqStore, parse_source, read_fasta, read_source  # noqa: F401

__all__ = ["AlignConfig", "Aligner", "MultiAligner", "align_all", "matrix_lookup", "matrix_names", "SeqStore",
           "ParseError", "parse_source", "read_fasta", "read_source"]
