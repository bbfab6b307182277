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


@pytest.mark.gpu
@pytest.mark.parametrize("name", golden_cases())
def test_dropin_matches_reference_cpu_build(name):
    from oracle import oracle
    ensure_built()
    g = load_golden(name)
    with tempfile.TemporaryDirectory() as td:
        fa = Path(td) / "in.fasta"
        fa.write_bytes(b"".join(b">s%d\n%s\n" % (k, s) for k, s in enumerate(g["seqs"])))
        res = oracle.run_reference(fa, g["args"], threads=1, binary=BIN)
    assert res["triangular"] == g["triangular"]
    assert res["seqs"] == g["kept"]
    assert np.array_equal(res["matrix"], g["matrix"])


@pytest.mark.gpu
def test_dropin_no_write_and_benchmark_readout():
    from oracle import oracle
    from sequencealigner_b200 import synth
    ensure_built()
    store, c = synth.make("c2", n=400)
    with tempfile.TemporaryDirectory() as td:
        fa = Path(td) / "in.fasta"
        store.to_fasta(fa)
        res = oracle.run_reference(fa, oracle.ref_args(**c), threads=0, write=False, bench=True,
                                   binary=BIN)
    assert res["align_seconds"] is not None and "Alignments per second" in res["stdout"]
