"""Synthetic workloads of BASELINE.json `configs` (definitions: SURVEY.md section 8d).

numpy `default_rng(seed)`, residues i.i.d. uniform over the stated alphabet.
`n` can be overridden to build scaled-down versions of the same distribution for
parity tests and the bounded CPU baseline.
"""
from __future__ import annotations

import numpy as np

from .seqstore import SeqStore

PROTEIN20 = "ARNDCQEGHILKMFPSTWYV"
DNA4 = "ACGT"


def _uniform(rng, n, lo, hi, alphabet):
    lens = rng.integers(lo, hi + 1, size=n)
    codes = rng.integers(0, len(alphabet), size=int(lens.sum()))
    return SeqStore.from_codes(codes, lens, alphabet)


def _with_duplicates(rng, n, lo, hi, alphabet, frac, max_subst):
    """A fraction of the sequences are copies of an earlier one with <= max_subst
    substitutions (C4: exercises the -f filter)."""
    seqs = []
    a = np.frombuffer(alphabet.encode(), dtype=np.uint8)
    for k in range(n):
        if k > 0 and rng.random() < frac:
            s = seqs[int(rng.integers(0, k))].copy()
            for _ in range(int(rng.integers(0, max_subst + 1))):
                s[int(rng.integers(0, s.shape[0]))] = a[int(rng.integers(0, a.shape[0]))]
        else:
            s = a[rng.integers(0, a.shape[0], size=int(rng.integers(lo, hi + 1)))]
        seqs.append(s)
    return SeqStore.from_sequences([s.tobytes() for s in seqs])


CONFIGS = {
    # id: (seed, n, lo, hi, alphabet, matrix, method, gaps, filter)
    "c1": dict(seed=1001, n=2000, lo=10, hi=60, alphabet=PROTEIN20, matrix="blosum62",
               method="nw", gap_pen=4),
    "c2": dict(seed=1002, n=10000, lo=250, hi=350, alphabet=PROTEIN20, matrix="blosum62",
               method="ga", gap_open=10, gap_ext=1),
    "c3": dict(seed=1003, n=50000, lo=150, hi=150, alphabet=DNA4, matrix="nuc44",
               method="sw", gap_open=10, gap_ext=1),
    "c4": dict(seed=1004, n=100000, lo=30, hi=60, alphabet=PROTEIN20, matrix="blosum62",
               method="ga", gap_open=10, gap_ext=1, filter=0.9, dup_frac=0.2, dup_subst=2),
    "c5": dict(seed=1005, n=200000, lo=50, hi=1000, alphabet=PROTEIN20, matrix="blosum62",
               method="sw", gap_open=10, gap_ext=1),
}


def cli_args(method: str, matrix: str, gap_pen=None, gap_open=None, gap_ext=None, filter=0.0,
             **_) -> list[str]:
    """The reference CLI's spelling of a workload (src/bio/align.c:111-202, filter.c:103-117)."""
    args = ["-m", matrix, "-a", method]
    if gap_pen is not None:
        args += ["-p", str(gap_pen)]
    if gap_open is not None:
        args += ["-s", str(gap_open)]
    if gap_ext is not None:
        args += ["-e", str(gap_ext)]
    if filter and filter > 0:
        args += ["-f", repr(float(filter))]
    return args


def make(config: str, n: int | None = None, seed: int | None = None) -> tuple[SeqStore, dict]:
    cfg = dict(CONFIGS[config])
    if n is not None:
        cfg["n"] = int(n)
    if seed is not None:
        cfg["seed"] = int(seed)
    rng = np.random.default_rng(cfg["seed"])
    if cfg.get("dup_frac"):
        store = _with_duplicates(rng, cfg["n"], cfg["lo"], cfg["hi"], cfg["alphabet"],
                                 cfg["dup_frac"], cfg["dup_subst"])
    else:
        store = _uniform(rng, cfg["n"], cfg["lo"], cfg["hi"], cfg["alphabet"])
    return store, cfg


def make_overflow(n: int = 2000, planted: int = 16, length: int = 1000,
                  seed: int = 1005) -> tuple[SeqStore, dict]:
    """C5-ovf: mixed lengths 50..1000 with planted identical poly-W sequences so that
    SW under pam500 (W:W = 34) exceeds int16 and the int32 re-run list is used."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(50, 1001, size=n)
    seqs = []
    a = np.frombuffer(PROTEIN20.encode(), dtype=np.uint8)
    where = set(rng.choice(n, size=planted, replace=False).tolist())
    for k in range(n):
        if k in where:
            seqs.append(b"W" * length)
        else:
            seqs.append(a[rng.integers(0, 20, size=int(lens[k]))].tobytes())
    cfg = dict(seed=seed, n=n, matrix="pam500", method="sw", gap_open=10, gap_ext=1)
    return SeqStore.from_sequences(seqs), cfg
