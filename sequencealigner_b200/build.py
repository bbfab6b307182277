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
SOURCES = ["sab_capi.cu", "sab_module.cu", "sab_filter.cu", "sab_tables.cpp", "sab_h5.cpp", "sab_source.cpp", "sab_noalloc.cu"]
# the strip kernel: one cubin per register-strip width (-DSAB_K=<columns>), embedded in the library and
# loaded by sab_module.cu, which overwrites the placeholder immediate below with the job's gap constant
WAVE_INST = "sab_wave_inst.cu"
LG_PLACEHOLDER = "0x5ab75ab7u"  # == sab::kLgPlaceholder (sab_kernels.cuh)
HEADERS = ["sab_kernels.cuh", "sab_wave.cuh", "sab_internal.h", "sab_module.h", "submat_tables.inc",
           "../../include/seqalign_b200.h"]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC,-O3,-Wall,-Wno-unused-function", "-cudart", "shared",
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


def strip_widths() -> list[int]:
    """The widths the kernel is instantiated for: the SAB_WIDTHS list of csrc/sab_kernels.cuh."""
    import re
    m = re.search(r"#define SAB_WIDTHS\(X\)(.*)", (CSRC / "sab_kernels.cuh").read_text())
    return [int(k) for k in re.findall(r"X\((\d+)\)", m.group(1))]


def build_library(force: bool = False, verbose: bool = False, out: Path | None = None,
                  defines: tuple[str, ...] = ()) -> Path:
    """nvcc -c every translation unit (in parallel: the strip kernel is ~64 KB of unrolled code per
    instantiation, 24 of them in 9 cubins), embed the cubins, then link.  `out` / `defines` build tuning variants."""
    from concurrent.futures import ThreadPoolExecutor
    out = Path(out or LIB)
    srcs = [CSRC / s for s in SOURCES]
    deps = srcs + [CSRC / WAVE_INST] + [CSRC / h for h in HEADERS]
    if not (force or _stale(out, deps)):
        return out
    objdir = PKG / "_obj" / out.stem
    objdir.mkdir(parents=True, exist_ok=True)
    base = [nvcc(), *NVCC_FLAGS, *defines] + (["-Xptxas=-v"] if verbose else [])
    jobs = [(base + ["-c", str(src), "-o", str(objdir / (src.stem + ".o"))]) for src in srcs]
    widths = strip_widths()
    cubin_flags = [f for f in NVCC_FLAGS if f not in ("-Xcompiler", "-cudart", "shared") and not f.startswith("-fPIC")]
    jobs += [([nvcc(), *cubin_flags, *defines] + (["-Xptxas=-v"] if verbose else []) +
              [f"-DSAB_K={k}", f"-DSAB_LG_IMM={LG_PLACEHOLDER}", "-cubin", str(CSRC / WAVE_INST),
               "-o", str(objdir / f"sab_wave_k{k}.cubin")]) for k in widths]

    def run(cmd):
        return subprocess.run(cmd, cwd=str(CSRC), capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=max(2, (os.cpu_count() or 4))) as ex:
        results = list(ex.map(run, jobs))
    for cmd, r in zip(jobs, results):
        if verbose or r.returncode:
            print(" ".join(cmd[-4:]), r.stdout, r.stderr, sep="\n")
        if r.returncode:
            raise subprocess.CalledProcessError(r.returncode, cmd, r.stdout, r.stderr)
    # the device images become read-only data of the library (symbols sab_cubin_k<K> / _end)
    asm = objdir / "sab_images.S"
    asm.write_text("\t.section .rodata\n" + "".join(
        f"\t.balign 64\n\t.global sab_cubin_k{k}\nsab_cubin_k{k}:\n\t.incbin \"{objdir / f'sab_wave_k{k}.cubin'}\"\n"
        f"\t.global sab_cubin_k{k}_end\nsab_cubin_k{k}_end:\n" for k in widths) +
        "\t.section .note.GNU-stack,\"\",@progbits\n")
    subprocess.run(["gcc", "-c", "-fPIC", str(asm), "-o", str(objdir / "sab_images.o")], check=True)
    objs = [c[-1] for c in jobs[:len(srcs)]] + [str(objdir / "sab_images.o")]
    subprocess.run([nvcc(), "-shared", "-cudart", "shared", "-gencode", "arch=compute_100a,code=sm_100a",
                    "-o", str(out), *objs, "-lz"], check=True, cwd=str(CSRC))
    return out


def build_pipe_peak(force: bool = False) -> Path:
    """Measurement infrastructure for bench.py only: integer-pipe peaks (tools/pipe_peak.cu, the
    roofline denominator) and the ceiling of the bare DP register block (tools/cell_peak.cu)."""
    tools = PKG.parent / "tools"
    srcs = [tools / "pipe_peak.cu", tools / "cell_peak.cu"]
    if force or _stale(PIPE_LIB, srcs + [CSRC / "sab_wave.cuh", CSRC / "sab_kernels.cuh"]):
        subprocess.run([nvcc(), *NVCC_FLAGS, "-shared", "-DCELL_PEAK_LIB", "-I", str(CSRC), "-I", str(PKG.parent / "include"),
                        "-o", str(PIPE_LIB), *map(str, srcs)], check=True)
    return PIPE_LIB


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 2 and sys.argv[1] == "--variant":  # tuning builds: --variant <name> [-D...]
        (PKG / "_variants").mkdir(exist_ok=True)
        print(build_library(force=True, out=PKG / "_variants" / f"{sys.argv[2]}.so",
                            defines=tuple(sys.argv[3:])))
    else:
        print(build_library(force=True, verbose="-v" in sys.argv))
        print(build_pipe_peak(force=True))
