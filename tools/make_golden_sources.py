#!/usr/bin/env python3
"""Generates tests/golden/sources.json: what the REFERENCE's own parsers (src/io/source/fasta.c,
src/io/source/dsv.c, reached through the reference CPU build oracle/_ref/seqalign) make of a set of
FASTA / delimiter-separated inputs -- the sequence list, or the one error line it prints.
tests/test_sources.py replays them through sab_parse_fasta / sab_parse_dsv.

    python tools/make_golden_sources.py          (needs oracle/_ref/seqalign: make -C oracle ref)
"""
import base64
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import oracle  # noqa: E402

# (name, extension, file bytes, options)
HAND = [
    # ---- FASTA, fasta.c:16-85 --------------------------------------------------------------
    ("fa_plain", "fasta", b">a\nACDEFG\n>b\nACDQFG\n>c\nWWWW\n", {}),
    ("fa_multiline_crlf_lower_blanks", "fa", b">x one\r\nac d\r\nEF\r\n\r\n>y\r\nww\r\n", {}),
    ("fa_no_final_newline", "fas", b">a\nAC\n>b\nDE", {}),
    ("fa_header_ends_with_cr_only", "fna", b">a\rAC\r>b\rDE\r", {"matrix": "nuc44"}),
    ("fa_blank_lines_after_header", "ffn", b">a\n\n\r\n\nAC\n>b\n\nGT\n", {"matrix": "nuc44"}),
    ("fa_gt_inside_a_data_line_starts_a_record", "faa", b">a\nAC>b\nDE\n", {}),
    ("fa_gt_inside_a_header_is_text", "frn", b">a >still header> x\nAC\n>b\nGU\n", {}),
    ("fa_upper_case_extension", "MPFA", b">a\nAC\n>b\nDE\n", {}),
    ("fa_stop_and_ambiguity_letters", "fasta", b">a\nBZX*\n>b\nacdefghiklmnpqrstvwy\n", {}),
    ("fa_letter_outside_blosum", "fasta", b">a\nACD\n>b\nAJD\n", {}),
    ("fa_protein_letter_under_nuc44", "fasta", b">a\nACGT\n>b\nACGE\n", {"matrix": "nuc44"}),
    ("fa_iupac_under_nuc44", "fasta", b">a\nacgtswrykmbvhdn*\n>b\nNNNN\n", {"matrix": "nuc44"}),
    ("fa_tab_is_not_blank", "fasta", b">a\nAC\tD\n>b\nAC\n", {}),
    ("fa_digit", "fasta", b">a\nAC\n>b\nA1C\n", {}),
    ("fa_nul_byte", "fasta", b">a\nAC\n>b\nA\0C\n", {}),
    ("fa_high_byte", "fasta", b">a\nAC\n>b\nA\xc3\xa9C\n", {}),
    ("fa_data_before_first_header", "fasta", b"AC\n>a\nAC\n>b\nAC\n", {}),
    ("fa_blank_line_before_first_header", "fasta", b"\n>a\nAC\n>b\nAC\n", {}),
    ("fa_last_header_without_data", "fasta", b">a\nAC\n>b\n", {}),
    ("fa_last_header_without_line_end", "fasta", b">a\nAC\n>b", {}),
    ("fa_empty_record_in_the_middle", "fasta", b">a\nAC\n>b\n>c\nAC\n", {}),
    ("fa_record_of_blanks_only", "fasta", b">a\nAC\n>b\n  \n \n>c\nAC\n", {}),
    ("fa_single_record", "fasta", b">a\nACDEF\n", {}),
    ("fa_length_limit_from_gap_penalty", "fasta", b">a\nAC\n>b\nACD\n", {"gap_pen": 500000000}),
    ("fa_length_limit_zero", "fasta", b">a\nA\n>b\nC\n", {"gap_pen": 2000000000}),
    ("fa_length_limit_met_exactly", "fasta", b">a\nAC\n>b\nCD\n", {"gap_pen": 500000000}),
    ("ext_unknown", "txt", b">a\nAC\n>b\nDE\n", {}),
    ("ext_fastq_is_not_fasta", "fastq", b">a\nAC\n>b\nDE\n", {}),
    # ---- DSV, dsv.c:26-230 ------------------------------------------------------------------
    ("csv_plain", "csv", b"id,sequence\n1,ACD\n2,WWY\n3,KLM\n", {}),
    ("csv_keyword_any_case_first_match_wins", "csv", b"Protein,SEQ,x\nACD,WW,1\nKLM,YY,2\n", {}),
    ("csv_keyword_quoted", "csv", b'id,"Peptide"\n1,ACD\n2,WWY\n', {}),
    ("csv_crlf_blank_lines_no_final_newline", "csv", b"seq,n\r\nACD,1\r\n\r\n\r\nWWY,2", {}),
    ("csv_quoted_sequence_blanks_lowercase", "csv", b'id,seq\n1,"ac d"\n2," w W y "\n', {}),
    ("csv_delimiter_inside_quotes", "csv", b'name,seq,note\n"a,b",ACD,"x,y"\n"c",WWY,z\n', {}),
    ("csv_line_end_inside_quotes", "csv", b'name,seq\n"two\nlines",ACD\nq,"WW\nY"\n', {}),
    ("csv_doubled_quote_inside_quotes", "csv", b'name,seq\n"say ""hi"", ok",ACD\nq,"W""W""Y"\n', {}),
    ("csv_quote_in_the_middle_of_a_sequence", "csv", b'id,seq\n1,AC"D"E\n2,WW\n', {}),
    ("csv_trailing_delimiter_is_tolerated", "csv", b"id,seq\n1,ACD,\n2,WWY,\n", {}),
    ("csv_too_many_columns", "csv", b"id,seq\n1,ACD\n2,WWY,extra\n", {}),
    ("csv_too_few_columns", "csv", b"id,seq,len\n1,ACD,3\n2,WWY\n", {}),
    ("csv_no_sequence_column", "csv", b"id,len,seq\n1,3,ACD\n2\n", {}),
    ("csv_empty_sequence_field", "csv", b"id,seq\n1,ACD\n2,\n", {}),
    ("csv_sequence_field_of_quotes_only", "csv", b'id,seq\n1,ACD\n2,""""\n', {}),
    ("csv_sequence_of_blanks_only", "csv", b'id,seq\n1,ACD\n2,"  "\n', {}),
    ("csv_invalid_letter", "csv", b"id,seq\n1,ACD\n2,AJD\n", {}),
    ("csv_nul_byte", "csv", b"id,seq\n1,ACD\n2,A\0D\n", {}),
    ("csv_high_byte", "csv", b"id,seq\n1,ACD\n2,A\xffD\n", {}),
    ("csv_first_row_has_empty_column", "csv", b"id,,seq\n1,x,ACD\n2,y,WWY\n", {}),
    ("csv_header_only", "csv", b"id,seq\n", {}),
    ("csv_one_row", "csv", b"id,seq\n1,ACD\n", {}),
    ("csv_length_limit", "csv", b"seq\nAC\nACD\n", {"gap_pen": 500000000}),
    ("tsv_plain", "tsv", b"id\tdna\n1\tACGT\n2\tGGCC\n", {"matrix": "nuc44"}),
    ("tsv_comma_is_not_a_delimiter", "tsv", b"id\tseq\n1,x\tACD\n2,y\tWWY\n", {}),
    ("ssv_plain", "ssv", b"rna;id\nACGU;1\nGGCC;2\n", {}),
    ("psv_plain_upper_case_extension", "PSV", b"id|chain|x\n1|ACD|q\n2|WWY|r\n", {}),
    ("csv_amino_keyword", "csv", b"amino\nACD\nWWY\n", {}),
    # the reference asks: which column, and is the first row a header at all (dsv.c:146-160)
    ("csv_ask_column_with_header", "csv", b"id,foo\n1,ACD\n2,WWY\n", {"column": 1}),
    ("csv_ask_column_zero_with_header", "csv", b"foo,id\nACD,1\nWWY,2\n", {"column": 0}),
    ("csv_ask_no_header_line", "csv", b"1,ACD\n2,WWY\n3,KLM\n", {"column": 1, "header": False}),
    ("csv_ask_no_header_line_letters_in_first_row", "csv", b"ACD,x\nWWY,y\n", {"column": 0, "header": False}),
    ("csv_ask_no_header_first_row_invalid", "csv", b"id,foo\n1,ACD\n2,WWY\n", {"column": 1, "header": False}),
    ("csv_ask_wrong_column", "csv", b"id,foo\n1,ACD\n2,WWY\n", {"column": 0}),
    ("csv_ask_no_header_but_empty_first_row_column", "csv", b"ACD,,x\nWWY,,y\n", {"column": 0, "header": False}),
]


def fuzz_cases(seed: int, count: int):
    """Seeded inputs: byte soups around the characters the parsers give a meaning to (a third), and
    well-formed files -- random line structure, case, blanks, quoting -- hit by at most two random byte
    edits (the rest), so that both the accepting and the rejecting branches are reached deep in a file."""
    rng = np.random.default_rng(seed)
    out = []
    fa_bytes = [b"A", b"C", b"D", b"w", b"y", b"J", b">", b"\n", b"\n", b"\r", b" ", b"\n>", b"\n>h\n", b"\0", b"\t"]
    fa_p = np.array([8, 8, 8, 4, 4, .3, 1.5, 4, 4, 1, 2, 3, 4, .15, .15])
    dsv_bytes = [b"A", b"C", b"D", b"w", b"J", b",", b",", b'"', b"\n", b"\n", b"\r\n", b" ", b";", b'""', b"1", b"\xfe"]
    dsv_p = np.array([8, 8, 8, 4, .3, 4, 4, 2.5, 4, 4, 1, 1.5, 1, 1, .3, .1])
    edits = [b">", b"\n", b"\r", b",", b'"', b" ", b"J", b"\0", b"\xe9", b"", b"", b"A", b";", b"\t"]
    letters = np.frombuffer(b"ACDEFGHIKLMNPQRSTVWYacdwyBZX*", dtype=np.uint8)

    def seq():
        return letters[rng.integers(0, letters.size, int(rng.integers(1, 14)))].tobytes()

    def mutate(data):
        for _ in range(int(rng.choice([0, 0, 1, 1, 2]))):
            at = int(rng.integers(0, len(data) + 1))
            data = data[:at] + edits[int(rng.integers(0, len(edits)))] + data[at + int(rng.integers(0, 2)):]
        return data

    def fasta():
        eol = b"\r\n" if rng.random() < .3 else b"\n"
        recs = []
        for r in range(int(rng.integers(1, 5))):
            s, lines = seq(), []
            while s:
                k = int(rng.integers(1, 8))
                lines.append((b" " if rng.random() < .1 else b"") + s[:k])
                s = s[k:]
            recs.append(b">r%d x" % r + eol + (eol * int(rng.integers(0, 2))) + eol.join(lines))
        return eol.join(recs) + (eol if rng.random() < .7 else b"")

    def dsv(delim):
        cols = int(rng.integers(1, 4))
        at = int(rng.integers(0, cols))
        r = rng.random()
        key = [b"seq", b"Sequence", b"PEPTIDE", b"chain"][int(rng.integers(0, 4))] if r < .6 else b"data"
        opt = {} if r < .6 else dict(column=at, header=bool(r < .8))

        def field(text, may_quote=True):
            q = rng.random()
            if may_quote and q < .15:
                return b'"' + text + b'"'
            if may_quote and q < .25:  # delimiter, doubled quote or line end inside quotes
                return b'"' + text[:1] + [delim, b'""', b"\n", b" "][int(rng.integers(0, 4))] + text[1:] + b'"'
            return text
        eol = b"\r\n" if rng.random() < .3 else b"\n"
        rows = []
        if opt.get("header", True):
            rows.append(delim.join(field(key, False) if c == at else b"c%d" % c for c in range(cols)))
        for _ in range(int(rng.integers(1, 5))):
            rows.append(delim.join(field(seq()) if c == at else field(b"x%d" % rng.integers(0, 99)) for c in range(cols)))
        return eol.join(rows) + (eol if rng.random() < .7 else b""), opt

    for k in range(count):
        n = int(rng.integers(4, 48))
        if k % 6 == 0:
            body = b"".join(fa_bytes[i] for i in rng.choice(len(fa_bytes), n, p=fa_p / fa_p.sum()))
            out.append((f"fuzz_fa_{k:03d}", "fasta", (b">h\n" if rng.random() < .9 else b"") + body, {}))
        elif k % 6 == 1:
            body = b"".join(dsv_bytes[i] for i in rng.choice(len(dsv_bytes), n, p=dsv_p / dsv_p.sum()))
            r = rng.random()
            if r < .6:
                out.append((f"fuzz_dsv_{k:03d}", "csv", (b"id,seq\n" if r < .3 else b"seq,id\n") + body, {}))
            else:  # the reference will ask
                out.append((f"fuzz_dsv_{k:03d}", "csv", body,
                            dict(column=int(rng.integers(0, 2)), header=bool(rng.random() < .5))))
        elif k % 2 == 0:
            out.append((f"fuzz_fa_{k:03d}", "fa", mutate(fasta()), {}))
        else:
            ext, delim = [("csv", b","), ("tsv", b"\t"), ("ssv", b";"), ("psv", b"|")][int(rng.integers(0, 4))]
            data, opt = dsv(delim)
            out.append((f"fuzz_dsv_{k:03d}", ext, mutate(data), opt))
    return out


def expected(ext, data, opt):
    r = oracle.reference_parse(data, ext, matrix=opt.get("matrix", "blosum62"), gap_pen=opt.get("gap_pen", 1),
                               column=opt.get("column"), header=opt.get("header", True))
    if opt.get("column") is not None and r["choices"] is not None and opt["column"] >= r["choices"] - 1:
        return None  # the answer does not exist for this file (fuzz): not a case
    if r["choices"] == 2 and not opt.get("header", True):
        # One column and "No header line": the reference's second prompt gets a 1-entry list, which clix
        # refuses (third_party/clix/print.h:1047-1051 returns -2), parse_dsv takes that as the column and
        # every such file dies with "DSV row #1 has too few columns".  Not reproduced (tests/test_sources.py).
        return None
    if r["seqs"] is not None:
        return dict(seqs=[s.decode("latin-1") for s in r["seqs"]], asked=r["choices"] is not None)
    return dict(error=r["error"], asked=r["choices"] is not None)


def main():
    assert oracle.ref_available(), "oracle/_ref/seqalign missing: make -C oracle ref"
    cases = []
    for name, ext, data, opt in HAND + fuzz_cases(20261105, 240):
        exp = expected(ext, data, opt)
        if exp is None:
            continue
        if "column" in opt and not exp["asked"]:
            opt = {k: v for k, v in opt.items() if k not in ("column", "header")}  # it found a header name
        elif exp["asked"] and "column" not in opt:
            opt = dict(opt, column=0, header=True)  # what reference_parse answers by default
        cases.append(dict(name=name, ext=ext, data=base64.b64encode(data).decode(), options=opt, expect=exp))
    out = ROOT / "tests" / "golden" / "sources.json"
    out.write_text(json.dumps(dict(
        generator="tools/make_golden_sources.py", binary="oracle/_ref/seqalign (unmodified reference CPU build)",
        cases=cases), indent=0) + "\n")
    errs = sum("error" in c["expect"] for c in cases)
    print(f"{out}: {len(cases)} cases ({errs} rejected inputs, {len(cases) - errs} accepted)")


if __name__ == "__main__":
    main()
