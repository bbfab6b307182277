#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the UNMODIFIED reference CPU build.

Run in the build container only (needs /root/reference, compiled into
oracle/_ref/seqalign by `make -C oracle`).  The reference ships no golden vectors
or tests of its own, so these fixtures -- outputs of the
reference's real main/parsers/filter/align path on small seeded inputs -- are what
pins the oracle restatement and the CUDA path on machines without the reference.

Each fixture: fasta (bytes), args (the CLI flags used), matrix (int32, layout as
written by the reference), triangular flag, kept (the post-filter sequence list).
"""
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import oracle  # noqa: E402

OUT = ROOT / "tests" / "golden"
P20 = "ARNDCQEGHILKMFPSTWYV"
P24 = "ARNDCQEGHILKMFPSTWYVBZX*"
N16 = "ATGCSWRYKMBVHDN*"


def rand_seqs(rng, n, lo, hi, alphabet):
    a = np.frombuffer(alphabet.encode(), dtype=np.uint8)
    return [a[rng.integers(0, len(a), size=int(rng.integers(lo, hi + 1)))].tobytes()
            for _ in range(n)]


def fasta_bytes(seqs):
    return b"".join(b">s%d\n%s\n" % (k, s) for k, s in enumerate(seqs))


def run(name, seqs, threads=1, **cli):
    fa = fasta_bytes(seqs)
    with tempfile.NamedTemporaryFile(suffix=".fasta") as f:
        f.write(fa)
        f.flush()
        args = oracle.ref_args(**cli)
        res = oracle.run_reference(f.name, args, threads=threads)
    np.savez_compressed(OUT / f"{name}.npz", fasta=np.frombuffer(fa, dtype=np.uint8),
                        args=np.array(args), matrix=res["matrix"],
                        triangular=np.array(res["triangular"]),
                        kept=np.array(res["seqs"], dtype=object).astype("S"))
    print(f"{name}: dim={len(res['seqs'])} tri={res['triangular']} "
          f"sum={int(res['matrix'].astype(np.int64).sum())}")


def extra():
    """Fixtures added after round 1; their own generator, so the ones above keep their bytes."""
    rng = np.random.default_rng(20261105)
    prot = rand_seqs(rng, 40, 1, 90, P20)
    # extend dearer than open (legal on the CLI, align.c:111-202): GA degenerates to NW with g = open
    run("protein_ga_blosum62_s2e5", prot, method="ga", matrix="blosum62", gap_open=2, gap_ext=5)
    run("protein_ga_blosum62_s0e1", prot, method="ga", matrix="blosum62", gap_open=0, gap_ext=1)
    run("protein_sw_blosum62_s2e5", prot, method="sw", matrix="blosum62", gap_open=2, gap_ext=5)
    run("protein_sw_blosum62_s0e1", prot, method="sw", matrix="blosum62", gap_open=0, gap_ext=1)
    iupac = rand_seqs(rng, 24, 10, 120, N16)
    run("iupac_ga_dnafull_s3e7", iupac, method="ga", matrix="dnafull", gap_open=3, gap_ext=7)


def main():
    assert oracle.ref_available(), "build oracle/_ref/seqalign first (make -C oracle)"
    OUT.mkdir(parents=True, exist_ok=True)
    if "--extra" in sys.argv:
        return extra()
    rng = np.random.default_rng(20261104)
    prot = rand_seqs(rng, 48, 1, 70, P20)
    run("protein_nw_blosum62_p4", prot, method="nw", matrix="blosum62", gap_pen=4)
    run("protein_ga_blosum62_s10e1", prot, method="ga", matrix="blosum62", gap_open=10, gap_ext=1)
    run("protein_sw_blosum62_s10e1", prot, method="sw", matrix="blosum62", gap_open=10, gap_ext=1)
    run("protein_ga_equal_s4e4", prot, method="ga", matrix="blosum62", gap_open=4, gap_ext=4)
    run("protein_nw_p0", prot[:24], method="nw", matrix="blosum62", gap_pen=0)
    run("protein_sw_s0e0", prot[:24], method="sw", matrix="blosum62", gap_open=0, gap_ext=0)
    amb = rand_seqs(rng, 40, 5, 90, P24)
    run("ambig_ga_pam250_s12e2", amb, method="ga", matrix="pam250", gap_open=12, gap_ext=2)
    run("ambig_sw_pam10_s20e3", amb, method="sw", matrix="pam10", gap_open=20, gap_ext=3)
    dna = rand_seqs(rng, 40, 150, 150, "ACGT")
    run("dna_sw_nuc44_s10e1", dna, method="sw", matrix="nuc44", gap_open=10, gap_ext=1)
    iupac = rand_seqs(rng, 32, 20, 200, N16)
    run("iupac_nw_dnafull_p5", iupac, method="nw", matrix="dnafull", gap_pen=5)
    run("iupac_ga_nuc44_s8e2", iupac, method="ga", matrix="nuc44", gap_open=8, gap_ext=2)
    mixed = rand_seqs(rng, 28, 50, 400, P20)
    run("mixed_sw_blosum45_s11e1", mixed, method="sw", matrix="blosum45", gap_open=11, gap_ext=1)
    run("mixed_ga_blosum80_s10e1", mixed, method="ga", matrix="blosum80", gap_open=10, gap_ext=1)
    # int16 overflow: identical poly-W under pam500 scores 34 per column
    ovf = rand_seqs(rng, 8, 200, 1000, P20) + [b"W" * 1000, b"W" * 990, b"W" * 1000,
                                               b"WY" * 500]
    run("overflow_sw_pam500_s10e1", ovf, method="sw", matrix="pam500", gap_open=10, gap_ext=1)
    run("overflow_ga_pam500_s10e1", ovf, method="ga", matrix="pam500", gap_open=10, gap_ext=1)
    run("bigpen_nw_blosum62_p40", ovf, method="nw", matrix="blosum62", gap_pen=40)
    # filter: planted near-duplicates, sequential (-T 1) semantics
    base = rand_seqs(rng, 60, 30, 60, P20)
    dup = []
    a = np.frombuffer(P20.encode(), dtype=np.uint8)
    for s in base:
        dup.append(s)
        if rng.random() < 0.5:
            t = bytearray(dup[int(rng.integers(0, len(dup)))])
            for _ in range(int(rng.integers(0, 4))):
                t[int(rng.integers(0, len(t)))] = int(a[int(rng.integers(0, 20))])
            dup.append(bytes(t))
    run("filter09_ga_blosum62", dup, method="ga", matrix="blosum62", gap_open=10, gap_ext=1,
        filter=0.9)
    run("filter05_nw_blosum62", dup, method="nw", matrix="blosum62", gap_pen=4, filter=0.5)
    # every table x every method on one small mixed set
    small_p = rand_seqs(rng, 14, 3, 40, P24)
    small_n = rand_seqs(rng, 14, 3, 40, N16)
    import re
    import subprocess
    lst = subprocess.run([str(oracle.REF_BIN), "-l"], capture_output=True, text=True).stdout
    names = [w for w in lst.split() if re.fullmatch(r"blosum\d+|pam\d+|dnafull|nuc44", w)]
    names = sorted(set(n.lower() for n in names))
    assert len(names) == 67, (len(names), lst)
    allm = {}
    for name in names:
        seqs = small_n if name in ("dnafull", "nuc44") else small_p
        with tempfile.NamedTemporaryFile(suffix=".fasta") as f:
            f.write(fasta_bytes(seqs))
            f.flush()
            for method, cli in (("nw", dict(gap_pen=3)), ("ga", dict(gap_open=9, gap_ext=2)),
                                ("sw", dict(gap_open=9, gap_ext=2))):
                res = oracle.run_reference(f.name, oracle.ref_args(method=method, matrix=name,
                                                                   **cli), threads=1)
                allm[f"{name}/{method}"] = res["matrix"]
    np.savez_compressed(OUT / "all_tables.npz", protein=np.frombuffer(fasta_bytes(small_p), np.uint8),
                        nucleo=np.frombuffer(fasta_bytes(small_n), np.uint8), **allm)
    print(f"all_tables: {len(allm)} matrices")
    extra()


if __name__ == "__main__":
    main()
