"""Builds libseqalign_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree.

`nvcc` cross-compiles without a GPU, so this runs in the build container; the
resulting .so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libseqalign_b200.so"
PIPE_LIB = PKG / "libsab_pipepeak.so"
SOURCES = ["sab_capi.cu", "sab_filter.cu", "sab_tables.cpp", "sab_h5.cpp"]
HEADERS = ["sab_kernels.cuh", "sab_wave.cuh", "sab_internal.h", "submat_tables.inc", "../../include/seqalign_b200.h"]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC,-O3,-Wall,-Wno-unused-function", "-shared", "-cudart", "shared",
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: libseqalign_b200.so cannot be built")


def _stale(out: Path, deps: list[Path]) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    srcs = [CSRC / s for s in SOURCES if (CSRC / s).exists()]
    deps = srcs + [CSRC / h for h in HEADERS]
    if force or _stale(LIB, deps):
        cmd = [nvcc(), *NVCC_FLAGS, "-o", str(LIB), *map(str, srcs), "-lz"]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.run(cmd, check=True, cwd=str(CSRC))
    return LIB


def build_pipe_peak(force: bool = False) -> Path:
    """Measurement infrastructure for bench.py only: integer-pipe peaks (tools/pipe_peak.cu, the
    roofline denominator) and the ceiling of the bare DP register block (tools/cell_peak.cu)."""
    tools = PKG.parent / "tools"
    srcs = [tools / "pipe_peak.cu", tools / "cell_peak.cu"]
    if force or _stale(PIPE_LIB, srcs + [CSRC / "sab_wave.cuh", CSRC / "sab_kernels.cuh"]):
        subprocess.run([nvcc(), *NVCC_FLAGS, "-DCELL_PEAK_LIB", "-I", str(CSRC), "-I", str(PKG.parent / "include"),
                        "-o", str(PIPE_LIB), *map(str, srcs)], check=True)
    return PIPE_LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
    print(build_pipe_peak(force=True))
