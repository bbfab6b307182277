"""TEST INFRASTRUCTURE ONLY: ctypes access to the CPU oracle.

 * liboracle.so   -- C restatement of the reference recurrences (seqalign_oracle.c)
 * _ref/seqalign  -- the UNMODIFIED reference CPU sources compiled in place by
                     oracle/Makefile (hdf5.c swapped for a raw dump writer)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "liboracle.so"
REF_BIN = HERE / "_ref" / "seqalign"


class OracleConfig(C.Structure):
    _fields_ = [("method", C.c_int32), ("gap_pen", C.c_int32), ("gap_open", C.c_int32),
                ("gap_ext", C.c_int32), ("seq_lut", C.c_int32 * 128),
                ("sub_mat", C.c_int32 * 576)]


def build(ref: bool = True) -> None:
    """make -C oracle (liboracle.so, and _ref/seqalign when /root/reference exists)."""
    subprocess.run(["make", "-C", str(HERE), "liboracle.so"] + (["ref"] if ref else []),
                   check=True, stdout=subprocess.DEVNULL)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB.exists():
            build(ref=False)
        l = C.CDLL(str(LIB))
        l.oracle_pair.restype = C.c_int32
        l.oracle_pair.argtypes = [C.POINTER(OracleConfig), C.c_void_p, C.c_int32, C.c_void_p,
                                  C.c_int32]
        l.oracle_align_all.restype = C.c_int64
        l.oracle_align_all.argtypes = [C.POINTER(OracleConfig), C.c_void_p, C.c_void_p, C.c_int32,
                                       C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
        l.oracle_pairs.restype = None
        l.oracle_pairs.argtypes = [C.POINTER(OracleConfig), C.c_void_p, C.c_void_p, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_void_p]
        l.oracle_filter.restype = C.c_int32
        l.oracle_filter.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_float, C.c_void_p]
        _lib = l
    return _lib


def make_config(resolved: dict) -> OracleConfig:
    """`resolved` = AlignConfig.resolve(): method id, gaps <= 0, seq_lut, sub_mat."""
    cfg = OracleConfig()
    cfg.method = int(resolved["method"])
    cfg.gap_pen, cfg.gap_open, cfg.gap_ext = (int(resolved["gap_pen"]), int(resolved["gap_open"]),
                                              int(resolved["gap_ext"]))
    cfg.seq_lut[:] = [int(v) for v in np.asarray(resolved["seq_lut"]).reshape(128)]
    cfg.sub_mat[:] = [int(v) for v in np.asarray(resolved["sub_mat"]).reshape(576)]
    return cfg


def align_all(resolved: dict, store, triangular: bool = True, first_j: int = 1,
              last_j: int | None = None, threads: int | None = None):
    """Returns (matrix, cells).  Columns j in [first_j, last_j) only (all by default)."""
    n = store.num
    out = np.zeros(n * (n - 1) // 2 if triangular else (n, n), dtype=np.int32)
    cfg = make_config(resolved)
    seqs = np.ascontiguousarray(store.seqs)
    meta = np.ascontiguousarray(store.meta)
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    cells = lib().oracle_align_all(C.byref(cfg), seqs.ctypes.data, meta.ctypes.data, n, store.max,
                                   out.ctypes.data, int(triangular), first_j,
                                   n if last_j is None else last_j)
    return out, int(cells)


def time_columns(resolved: dict, store, first_j: int, last_j: int):
    """Scores columns [first_j, last_j) without storing (CPU baseline timing)."""
    cfg = make_config(resolved)
    seqs = np.ascontiguousarray(store.seqs)
    meta = np.ascontiguousarray(store.meta)
    return int(lib().oracle_align_all(C.byref(cfg), seqs.ctypes.data, meta.ctypes.data, store.num,
                                      store.max, None, 1, first_j, last_j))


def pair(resolved: dict, seq_j: bytes, seq_i: bytes) -> int:
    cfg = make_config(resolved)
    return int(lib().oracle_pair(C.byref(cfg), seq_j, len(seq_j), seq_i, len(seq_i)))


def pairs(resolved: dict, store, i: np.ndarray, j: np.ndarray) -> np.ndarray:
    """Scores the listed (i<j) pairs (all host threads)."""
    cfg = make_config(resolved)
    seqs = np.ascontiguousarray(store.seqs)
    meta = np.ascontiguousarray(store.meta)
    i = np.ascontiguousarray(i, dtype=np.int32)
    j = np.ascontiguousarray(j, dtype=np.int32)
    out = np.zeros(i.size, dtype=np.int32)
    lib().oracle_pairs(C.byref(cfg), seqs.ctypes.data, meta.ctypes.data, i.size, i.ctypes.data,
                       j.ctypes.data, out.ctypes.data)
    return out


def filter_lost(store, threshold: float) -> np.ndarray:
    lost = np.zeros(store.num, dtype=np.uint8)
    seqs = np.ascontiguousarray(store.seqs)
    meta = np.ascontiguousarray(store.meta)
    lib().oracle_filter(seqs.ctypes.data, meta.ctypes.data, store.num, C.c_float(threshold),
                        lost.ctypes.data)
    return lost.astype(bool)


# --- the real reference binary ----------------------------------------------------------

def ref_binary() -> Path:
    """The reference build for this host: the reference compiles with -march=native
    (CMakeLists.txt:7-9); the prebuilt binaries are x86-64-v3 and -v4, take the highest level
    this CPU runs."""
    v4 = REF_BIN.with_name("seqalign_v4")
    if v4.exists():
        try:
            flags = next((l for l in open("/proc/cpuinfo") if l.startswith("flags")), "")
            if all(f" {f}" in flags for f in ("avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl")):
                return v4
        except OSError:
            pass
    return REF_BIN


def ref_available() -> bool:
    return REF_BIN.exists() and os.access(REF_BIN, os.X_OK)


def read_raw_dump(path):
    """Format written by oracle/ref_flush_raw.c."""
    data = Path(path).read_bytes()
    magic, dim, tri = np.frombuffer(data[:24], dtype="<u8")
    assert int(magic) == 0x315741524C415153, "bad raw dump"
    dim, tri = int(dim), int(tri)
    n = dim * (dim - 1) // 2 if tri else dim * dim
    mat = np.frombuffer(data[24:24 + 4 * n], dtype="<i4").copy()
    if not tri:
        mat = mat.reshape(dim, dim)
    seqs = data[24 + 4 * n:].split(b"\0")[:dim]
    return mat, bool(tri), seqs


def ref_args(*a, **cli):
    from sequencealigner_b200.synth import cli_args  # the flag spelling lives with the workloads
    return cli_args(*a, **cli)


def run_reference(fasta_path, cli_args, threads: int = 0, write: bool = True, bench: bool = False,
                  binary: Path | None = None, env: dict | None = None):
    """Runs a `seqalign` binary (the reference CPU build by default) on a FASTA file.
    Returns dict(matrix, triangular, seqs, align_seconds, stdout)."""
    binary = Path(binary or ref_binary())
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "out.bin"
        # `-T 0` resolves through omp_get_max_threads() (src/system/os.c:468-473), which obeys
        # OMP_NUM_THREADS -- torchrun exports OMP_NUM_THREADS=1 -- so "all host threads" is
        # passed as an explicit count and the OMP_* variables are not inherited.
        nthreads = threads if threads > 0 else (len(os.sched_getaffinity(0)) or os.cpu_count() or 1)
        cmd = [str(binary), "-i", str(fasta_path), *cli_args, "-F", "-P", "-T", str(nthreads)]
        cmd += ["-o", str(out)] if write else ["-W"]
        cmd += ["-B"] if bench else ["-Q"]
        e = {k: v for k, v in os.environ.items() if not k.startswith(("OMP_", "GOMP_", "KMP_"))}
        e.update(env or {})
        p = subprocess.run(cmd, capture_output=True, text=True, env=e)
        if p.returncode != 0:
            raise RuntimeError(f"reference binary failed ({p.returncode}): {p.stdout}\n{p.stderr}")
        res = dict(stdout=p.stdout, matrix=None, triangular=None, seqs=None, align_seconds=None,
                   threads=None, filter_seconds=None, binary=binary.name)
        m = re.search(r"CPU Threads:\s*(\d+)", p.stdout)  # what the reference itself reports
        if m:
            res["threads"] = int(m.group(1))
        m = re.search(r"Filtering:\s*([0-9.]+)\s*sec", p.stdout)
        if m:
            res["filter_seconds"] = float(m.group(1))
        if write:
            res["matrix"], res["triangular"], res["seqs"] = read_raw_dump(out)
        if bench:
            m = re.search(r"Alignment:\s*([0-9.]+)\s*sec", p.stdout)
            if m:
                res["align_seconds"] = float(m.group(1))
        return res


def reference_parse(data: bytes, ext: str, matrix: str = "blosum62", gap_pen: int = 1,
                    column: int | None = None, header: bool = True, binary: Path | None = None,
                    extra_args=(), read_sequences=None):
    """What the reference's input_load makes of a file (src/io/input.c:28-81 with the parsers
    src/io/source/{fasta,dsv}.c): runs a `seqalign` binary on it and returns
    dict(seqs=[...] | None, error=<the perr line> | None, choices=<entries of the column prompt>).
    A DSV file without a recognised header name makes the reference ASK (dsv.c:146-160); the
    answers are `column` (zero-based) and `header` (False = its "No header line" entry).
    `binary` / `extra_args` / `read_sequences(path)`: the same run with another build of the CLI
    (the drop-in, whose output file is HDF5)."""
    binary = Path(binary or ref_binary())
    with tempfile.TemporaryDirectory() as td:
        src = Path(td) / f"in.{ext}"
        src.write_bytes(data)
        out = Path(td) / "out.bin"
        cmd = [str(binary), "-i", str(src), "-o", str(out), "-m", matrix, "-a", "nw", "-p",
               str(gap_pen), "-T", "1", "-Q", "-P", "-F", *extra_args]
        e = {k: v for k, v in os.environ.items() if not k.startswith(("OMP_", "GOMP_", "KMP_"))}

        def run(answers):
            # the prompt re-asks forever at end of input: pad with an always-valid answer
            feed = "".join(f"{a}\n" for a in answers) + "1\n" * 8
            return subprocess.run(cmd, input=feed, capture_output=True, text=True, env=e, timeout=60,
                                  errors="replace")
        p = run([])
        choices = None
        m = re.search(r"Enter column number \(1-(\d+)\)", p.stdout)
        if m:  # it asked: run again with the caller's answers
            choices = int(m.group(1))
            if out.exists():
                out.unlink()
            col = 0 if column is None else column
            p = run([col + 1] if header else [choices, col + 1])
        if p.returncode == 0:
            seqs = read_sequences(out) if read_sequences else read_raw_dump(out)[2]
            return dict(seqs=seqs, error=None, choices=choices)
        lines = [l.strip() for l in p.stderr.splitlines() if l.strip()]
        return dict(seqs=None, error=lines[0] if lines else f"exit {p.returncode}", choices=choices)
