"""The reference CLI running on libseqalign_b200.so (integration/_build/seqalign_b200:
unmodified reference host sources + integration/seqalign_shim.c exporting cuda_align /
cuda_memory with the reference's signatures, src/interface/seqalign_cuda.h:7-9).

GPU tests replay the golden fixtures -- outputs of the reference's own CPU build for
the same FASTA and flags -- through that binary and compare the dumped matrix and
the post-filter sequence list bit for bit."""
import os
import subprocess
import tempfile
from pathlib import Path

import numpy as np
import pytest

from conftest import ROOT, golden_cases, load_golden

BIN = ROOT / "integration" / "_build" / "seqalign_b200"


def ensure_built():
    if Path("/root/reference/src").is_dir():
        from sequencealigner_b200.build import build_library
        build_library()
        subprocess.run(["make", "-C", str(ROOT / "integration")], check=True,
                       stdout=subprocess.DEVNULL)
    if not BIN.exists():
        pytest.skip("integration/_build/seqalign_b200 not built (needs the reference tree)")


def test_dropin_binary_links_the_library_and_keeps_the_cli():
    ensure_built()
    out = subprocess.run(["ldd", str(BIN)], capture_output=True, text=True).stdout
    assert "libseqalign_b200.so" in out
    hlp = subprocess.run([str(BIN), "-h"], capture_output=True, text=True).stdout
    for flag in ("-a, --align", "-m, --matrix", "-p, --gap-penalty", "-s, --gap-open",
                 "-e, --gap-extend", "-f, --filter", "-z, --compression", "-C, --no-cuda",
                 "-W, --no-write", "-B, --benchmark"):
        assert flag in hlp, flag
    syms = subprocess.run(["nm", "-D", "--defined-only", str(BIN)], capture_output=True,
                          text=True).stdout + \
        subprocess.run(["nm", str(BIN)], capture_output=True, text=True).stdout
    assert " cuda_align" in syms and " cuda_memory" in syms


def test_dropin_without_device_fails_loudly_instead_of_falling_back():
    ensure_built()
    import ctypes
    from sequencealigner_b200 import capi
    if capi.load().sab_device_count() > 0:
        pytest.skip("a device is present")
    with tempfile.TemporaryDirectory() as td:
        fa = Path(td) / "x.fasta"
        fa.write_bytes(b">a\nACDEFG\n>b\nACDQFG\n>c\nWWWW\n")
        p = subprocess.run([str(BIN), "-i", str(fa), "-o", str(Path(td) / "o.bin"), "-m",
                            "blosum62", "-a", "nw", "-p", "4", "-F", "-P"],
                           capture_output=True, text=True)
    assert p.returncode != 0 and not (Path(td) / "o.bin").exists()


def run_dropin(fasta, cli, extra=()):
    """Runs the drop-in CLI, returns (sequences, full matrix, stdout) read back from its HDF5."""
    import h5min
    out = Path(fasta).with_suffix(".h5")
    cmd = [str(BIN), "-i", str(fasta), "-o", str(out), *cli, "-F", "-P", *extra]
    p = subprocess.run(cmd, capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    seqs, mat, info = h5min.read_similarity_file(out)
    return seqs, mat, p.stdout


def golden_full(g):
    m = g["matrix"]
    if not g["triangular"]:
        return m
    n = len(g["kept"])
    full = np.zeros((n, n), dtype=np.int32)
    full[np.tril_indices(n, -1)] = m
    return full + full.T


def write_fasta(td, seqs):
    fa = Path(td) / "in.fasta"
    fa.write_bytes(b"".join(b">s%d\n%s\n" % (k, s) for k, s in enumerate(seqs)))
    return fa


@pytest.mark.parametrize("name", ["protein_ga_blosum62_s10e1", "filter09_ga_blosum62",
                                  "dna_sw_nuc44_s10e1"])
def test_dropin_cli_and_hdf5_writer_on_the_reference_cpu_path(name):
    """No GPU needed: with -C the reference's own CPU align() fills the matrix, so this pins the
    rest of the drop-in -- CLI, filter, output_load and the library's HDF5 writer -- to the
    golden output of the reference's CPU build."""
    ensure_built()
    g = load_golden(name)
    with tempfile.TemporaryDirectory() as td:
        seqs, mat, _ = run_dropin(write_fasta(td, g["seqs"]), g["args"], ["-C", "-T", "1", "-z", "4"])
    assert seqs == g["kept"]
    assert np.array_equal(mat, golden_full(g))


def test_dropin_parsers_on_the_reference_parser_goldens(oracle_mod):
    """No GPU needed (-C): the drop-in's FASTA / DSV plugins (integration/source_b200.c on
    sab_parse_fasta / sab_parse_dsv), prompts included, against what the reference's own
    src/io/source/{fasta,dsv}.c made of the same files (tests/golden/sources.json)."""
    import base64
    import json
    import h5min
    ensure_built()
    cases = json.loads((ROOT / "tests" / "golden" / "sources.json").read_text())["cases"]
    assert len(cases) > 250
    for c in cases:
        o = c["options"]
        r = oracle_mod.reference_parse(base64.b64decode(c["data"]), c["ext"], matrix=o.get("matrix", "blosum62"),
                                       gap_pen=o.get("gap_pen", 1), column=o.get("column"),
                                       header=o.get("header", True), binary=BIN, extra_args=["-C"],
                                       read_sequences=lambda f: h5min.read_similarity_file(f)[0])
        want = c["expect"]
        assert (r["choices"] is not None) == want["asked"], (c["name"], r)
        if "seqs" in want:
            assert r["seqs"] == [s.encode("latin-1") for s in want["seqs"]], (c["name"], r, want)
        else:
            assert r["error"] == want["error"], (c["name"], r, want)
    # the one-column file answered "No header line", which the reference cannot get past (its second
    # prompt is refused, tests/test_sources.py): read as that column
    r = oracle_mod.reference_parse(b"ACD\nWWY\nKLM", "csv", column=0, header=False, binary=BIN, extra_args=["-C"],
                                   read_sequences=lambda f: h5min.read_similarity_file(f)[0])
    assert r["seqs"] == [b"ACD", b"WWY", b"KLM"] and r["choices"] == 2


@pytest.mark.gpu
@pytest.mark.parametrize("name", golden_cases())
def test_dropin_matches_reference_cpu_build(name):
    ensure_built()
    g = load_golden(name)
    with tempfile.TemporaryDirectory() as td:
        seqs, mat, out = run_dropin(write_fasta(td, g["seqs"]), g["args"], ["-T", "1"])
    assert "CUDA: Enabled (libseqalign_b200" in out
    assert seqs == g["kept"]
    assert np.array_equal(mat, golden_full(g))


@pytest.mark.gpu
def test_dropin_chunked_compressed_output_and_benchmark_readout(oracle_mod):
    from sequencealigner_b200 import AlignConfig, synth
    ensure_built()
    store, c = synth.make("c2", n=600)
    cfg = AlignConfig(c["method"], c["matrix"], gap_open=c["gap_open"], gap_ext=c["gap_ext"])
    want, _ = oracle_mod.align_all(cfg.resolve(), store, triangular=False)
    with tempfile.TemporaryDirectory() as td:
        fa = Path(td) / "in.fasta"
        store.to_fasta(fa)
        seqs, mat, out = run_dropin(fa, oracle_mod.ref_args(**c), ["-z", "3", "-B", "-T", "0"])
        p = subprocess.run([str(BIN), "-i", str(fa), *oracle_mod.ref_args(**c), "-F", "-P", "-W",
                            "-B"], capture_output=True, text=True)
    assert seqs == [store.sequence(k) for k in range(store.num)]
    assert np.array_equal(mat, want)
    assert "Alignments per second" in out
    assert p.returncode == 0 and "Alignments per second" in p.stdout  # -W: no output file
