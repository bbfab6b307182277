"""CPU-only: the input formats in front of the path.  sab_parse_fasta / sab_parse_dsv
(csrc/sab_source.cpp, through the Python host mirror `parse_source`) against the reference's own
parsers (src/io/source/fasta.c:16-85, src/io/source/dsv.c:26-230, src/io/input.c:28-81):

 * tests/golden/sources.json -- outputs of the reference CPU build for hand-written and seeded
   byte-soup inputs (tools/make_golden_sources.py): the sequence list, or the error line;
 * live, where oracle/_ref/seqalign exists: fresh byte soups run through both.
"""
import base64
import collections
import json

import numpy as np
import pytest

from conftest import ROOT
from sequencealigner_b200 import ParseError, capi, matrix_lookup, parse_source, read_source

GOLDEN = json.loads((ROOT / "tests" / "golden" / "sources.json").read_text())["cases"]


def ours(ext, data, opt):
    """-> ("seqs", [...]) or ("error", message), like the fixture's `expect`."""
    lut, _ = matrix_lookup(opt.get("matrix", "blosum62"))
    try:
        store = parse_source(data, ext, lut, gap_pen=-opt.get("gap_pen", 1), column=opt.get("column"),
                             header=opt.get("header", True))
    except ParseError as e:
        return "error", str(e)
    assert store.meta["off"].tolist() == np.concatenate(
        [[0], np.cumsum(store.meta["len"].astype(np.int64) + 1)[:-1]]).tolist()
    assert store.seqs.size == int((store.meta["len"].astype(np.int64) + 1).sum())
    return "seqs", [store.sequence(k).decode("latin-1") for k in range(store.num)]


def same_error(got: str, want: str) -> bool:
    # the reference names the file in this one message (src/io/input.c:60)
    if want.startswith("Unsupported file format:"):
        return got.startswith("Unsupported file format:")
    return got == want


def test_golden_covers_every_message_and_both_formats():
    msgs = collections.Counter()
    for c in GOLDEN:
        e = c["expect"].get("error")
        if e:
            msgs[" ".join(w for w in e.split() if not w.lstrip("#").isdigit())] += 1
    for m in ("Data before first header", "Last header has no data", "Sequence is empty", "Sequence is invalid",
              "Sequence is corrupted", "Sequence exceeds length limits", "First row has empty column",
              "DSV row has no sequence column", "DSV row has too few columns", "DSV row has too many columns",
              "Not enough sequences: (min: 2)"):
        assert msgs[m] > 0, (m, msgs)
    ok = [c for c in GOLDEN if "seqs" in c["expect"]]
    assert sum(c["ext"].lower() in ("csv", "tsv", "ssv", "psv") for c in ok) >= 20
    assert sum(c["ext"].lower() not in ("csv", "tsv", "ssv", "psv") for c in ok) >= 20
    assert sum(c["expect"]["asked"] for c in GOLDEN) >= 10


@pytest.mark.parametrize("case", GOLDEN, ids=[c["name"] for c in GOLDEN])
def test_parsers_match_the_reference_golden(case):
    kind, got = ours(case["ext"], base64.b64decode(case["data"]), case["options"])
    want = case["expect"]
    if "seqs" in want:
        assert kind == "seqs" and got == want["seqs"], (kind, got, want)
    else:
        assert kind == "error" and same_error(got, want["error"]), (kind, got, want)


def test_a_file_without_a_sequence_column_name_needs_the_callers_answer():
    lut, _ = matrix_lookup("blosum62")
    with pytest.raises(ParseError, match="column="):
        parse_source(b"id,foo\n1,ACD\n2,WWY\n", "csv", lut)
    s = parse_source(b"id,foo\n1,ACD\n2,WWY\n", "csv", lut, column=1)
    assert [s.sequence(k) for k in range(s.num)] == [b"ACD", b"WWY"]
    with pytest.raises(ParseError, match="out of range"):
        parse_source(b"id,foo\n1,ACD\n2,WWY\n", "csv", lut, column=2)


def test_headerless_single_column_file_is_parsed_where_the_reference_cannot(oracle_mod):
    """Deliberate difference: for a one-column file answered "No header line" the reference's second
    prompt is a 1-entry list, which clix refuses with -2 (third_party/clix/print.h:1047-1051); parse_dsv
    uses that as the column (dsv.c:153) and every such file fails with 'too few columns'.  The library is
    told the column and parses it -- including the last row without a line end, whose NUL lands on
    file[bytes] (see the header)."""
    lut, _ = matrix_lookup("blosum62")
    s = parse_source(b"ACD\nWWY", "csv", lut, column=0, header=False)
    assert [s.sequence(k) for k in range(s.num)] == [b"ACD", b"WWY"]
    if oracle_mod.ref_available():
        r = oracle_mod.reference_parse(b"ACD\nWWY\n", "csv", column=0, header=False)
        assert r["error"] == "DSV row #1 has too few columns"


def test_read_source_picks_the_parser_from_the_extension(tmp_path):
    lut, _ = matrix_lookup("blosum62")
    (tmp_path / "a.b.FaA").write_bytes(b">x\nac\n>y\nde\n")
    (tmp_path / "t.tsv").write_bytes(b"n\tchain\n1\tAC\n2\tDE\n")
    for f in ("a.b.FaA", "t.tsv"):
        s = read_source(tmp_path / f, lut)
        assert [s.sequence(k) for k in range(s.num)] == [b"AC", b"DE"]
    (tmp_path / "noext").write_bytes(b">x\nAC\n>y\nDE\n")
    (tmp_path / ".fasta").write_bytes(b">x\nAC\n>y\nDE\n")
    for f in ("noext", ".fasta"):  # src/io/input.c:33-36
        with pytest.raises(ParseError, match="File extension not found"):
            read_source(tmp_path / f, lut)


def test_c_abi_reports_names_and_rejects_bad_arguments():
    import ctypes as C
    lib = capi.load()
    d = C.c_uint8(0)
    assert lib.sab_source_format(b"TSV", C.byref(d)) == capi.SAB_SRC_DSV and d.value == 9
    assert lib.sab_source_format(b"mpfa", None) == capi.SAB_SRC_FASTA
    assert lib.sab_source_format(b"fastq", None) == capi.SAB_SRC_NONE
    data = b'id,"My seq",Chain\n'
    buf = C.create_string_buffer(data, len(data))
    cols, found = C.c_int32(), C.c_int32()
    names = C.create_string_buffer(64)
    assert lib.sab_dsv_header(buf, len(data), ord(","), C.byref(cols), C.byref(found), names, 64, None, 0) == 0
    assert (cols.value, found.value) == (3, 2) and names.raw.split(b"\0")[:3] == [b"id", b"My seq", b"Chain"]
    assert lib.sab_dsv_header(buf, len(data), ord(","), C.byref(cols), C.byref(found), names, 4, None, 0) == -1
    lut = (C.c_int32 * 128)(*matrix_lookup("blosum62")[0].tolist())
    n, m = C.c_int32(), C.c_int32()
    fa = C.create_string_buffer(b">a\nAC\n>b\nD\n", 11)
    assert lib.sab_parse_fasta(fa, 11, lut, 1, C.byref(n), C.byref(m), None, 0) == -1  # gap_pen > 0
    assert lib.sab_parse_fasta(fa, 11, lut, 0, C.byref(n), C.byref(m), None, 0) == 0
    assert (n.value, m.value) == (2, 2) and fa.raw[:5] == b"AC\0D\0"
    meta = (capi.SabMeta * 2)()
    assert lib.sab_build_meta(fa, 2, meta) == 5
    assert [(x.off, x.len) for x in meta] == [(0, 2), (3, 1)]
    err = C.create_string_buffer(8)  # a short message buffer is filled, not overrun
    bad = C.create_string_buffer(b">a\nAJ\n", 6)
    assert lib.sab_parse_fasta(bad, 6, lut, 0, C.byref(n), C.byref(m), err, 8) == capi.SAB_ERR_PARSE
    assert err.value == b"Sequenc"


def test_large_fasta_round_trip_and_rate():
    """A C2-sized file (10 000 proteins, 60 letters per line): the store equals the generator's."""
    import time
    from sequencealigner_b200 import synth
    store, _ = synth.make("c2")
    lines = []
    for k in range(store.num):
        s = store.sequence(k)
        lines.append(b">s%d some description\n" % k + b"\n".join(s[i:i + 60] for i in range(0, len(s), 60)) + b"\n")
    data = b"".join(lines)
    lut, _ = matrix_lookup("blosum62")
    t0 = time.perf_counter()
    got = parse_source(data, "fasta", lut)
    dt = time.perf_counter() - t0
    assert np.array_equal(got.seqs, store.seqs) and np.array_equal(got.meta, store.meta)
    print(f"parse_source: {len(data) / 1e6:.1f} MB in {dt * 1e3:.1f} ms")


def test_live_byte_soups_against_the_reference_binary(oracle_mod):
    """Fresh inputs every run of the suite would hide regressions; the seed is fixed, but differs
    from the fixture's, so this doubles the sample where the reference binary is present."""
    if not oracle_mod.ref_available():
        pytest.skip("oracle/_ref/seqalign not built")
    import sys
    sys.path.insert(0, str(ROOT / "tools"))
    from make_golden_sources import fuzz_cases
    checked = 0
    for name, ext, data, opt in fuzz_cases(7, 120):
        r = oracle_mod.reference_parse(data, ext, column=opt.get("column"), header=opt.get("header", True))
        if r["choices"] is None:
            opt = {}
        elif "column" not in opt:
            opt = dict(column=0, header=True)  # reference_parse's default answers
        elif opt.get("column", 0) >= r["choices"] - 1 or (r["choices"] == 2 and not opt.get("header", True)):
            continue
        kind, got = ours(ext, data, opt)
        if r["seqs"] is not None:
            assert kind == "seqs" and got == [s.decode("latin-1") for s in r["seqs"]], (name, data, got, r)
        else:
            assert kind == "error" and same_error(got, r["error"]), (name, data, got, r)
        checked += 1
    assert checked >= 100


def test_live_long_lines_and_mixed_line_ends_against_the_reference_binary(oracle_mod):
    """The FASTA block-copy path: records of up to 400 letters in lines of 1..120, LF / CRLF / CR-only line
    ends, lower case, then up to three byte edits -- both parsers must agree on every file."""
    if not oracle_mod.ref_available():
        pytest.skip("oracle/_ref/seqalign not built")
    rng = np.random.default_rng(77)
    edits = [b">", b"\n", b"\r", b"\r\n", b" ", b"J", b"\0", b"\xe9", b"", b"", b"a", b"\t", b"\n\n", b"\r\r", b">x"]
    for matrix, alpha in (("blosum62", b"ACDEFGHIKLMNPQRSTVWYacdw"), ("nuc44", b"ACGTacgtNn")):
        letters = np.frombuffer(alpha, np.uint8)
        for _ in range(120):
            eol = [b"\n", b"\r\n", b"\r"][int(rng.integers(0, 3))] if rng.random() < .5 else b"\n"
            recs = []
            for r in range(int(rng.integers(2, 5))):
                s = letters[rng.integers(0, letters.size, int(rng.integers(1, 400)))].tobytes()
                w = int(rng.integers(1, 120))
                recs.append(b">r%d" % r + eol + eol.join(s[i:i + w] for i in range(0, len(s), w)))
            data = eol.join(recs) + (eol if rng.random() < .6 else b"")
            for _ in range(int(rng.choice([0, 1, 1, 2, 3]))):
                at = int(rng.integers(0, len(data) + 1))
                data = data[:at] + edits[int(rng.integers(0, len(edits)))] + data[at + int(rng.integers(0, 2)):]
            r = oracle_mod.reference_parse(data, "fa", matrix=matrix)
            kind, got = ours("fa", data, {"matrix": matrix})
            if r["seqs"] is not None:
                assert kind == "seqs" and got == [s.decode("latin-1") for s in r["seqs"]], (data, got, r)
            else:
                assert kind == "error" and got == r["error"], (data, got, r)
