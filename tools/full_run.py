#!/usr/bin/env python3
"""One BASELINE.json config at its stated size through the in-process multi-GPU path
(sab_multi_*: what the drop-in CLI's cuda_align drives), into ONE caller-owned matrix that may be
a file-backed MAP_SHARED mapping like the reference's tmp-file sink (src/io/output.c:35-55,
alloc_mmap src/system/os.c:103-140).  Prints one JSON line; saves a sample of the scores
(>= 1e6 uniformly drawn pairs + whole rows) to an .npz that tools/verify_samples.py re-scores with
the CPU oracle later, off the GPU clock.

    python tools/full_run.py --config c5 --ndev 8 --sink file --ordered --samples gpurun_out/c5.npz
"""
from __future__ import annotations

import argparse
import json
import mmap
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from sequencealigner_b200 import AlignConfig, MultiAligner, SeqStore, capi, synth  # noqa: E402


def mem_available() -> int:
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable:"):
            return int(line.split()[1]) * 1024
    return 0


def open_sink(kind: str, nbytes: int, directory: str | None):
    """('ram' | 'file' | 'auto') -> (int32 view, closer, description).  auto = the reference's rule:
    a tmp file when the FULL matrix would not fit in 3/4 of MemAvailable (output.c:36)."""
    if kind == "ram":
        return np.zeros(nbytes // 4, dtype=np.int32), (lambda: 0.0), "anonymous memory"
    # the reference uses /tmp (os.c:111); when its file system cannot hold the matrix, the next
    # candidate with room (a tmpfs mapping is still a file-backed MAP_SHARED sink, without write-back)
    import shutil
    cands = [directory] if directory else ["/tmp", "/var/tmp", os.getcwd(), "/dev/shm"]
    d = next((c for c in cands if c and os.path.isdir(c) and shutil.disk_usage(c).free > nbytes * 1.05), None)
    if d is None:
        raise SystemExit(f"no directory with {nbytes / 2**30:.0f} GiB free for the sink file: {cands}")
    fd = os.open(d, os.O_TMPFILE | os.O_RDWR, 0o600)  # os.c:111: unnamed file in /tmp
    os.ftruncate(fd, nbytes)
    mm = mmap.mmap(fd, nbytes, flags=mmap.MAP_SHARED)
    try:
        mm.madvise(mmap.MADV_RANDOM)  # os.c:128-131
        mm.madvise(mmap.MADV_HUGEPAGE)
    except (OSError, AttributeError, ValueError):
        pass
    view = np.frombuffer(mm, dtype=np.int32)

    def close():
        t0 = time.perf_counter()
        mm.flush()  # msync: what is still dirty goes to the file now
        dt = time.perf_counter() - t0
        os.close(fd)
        return dt
    real = os.path.realpath(d)
    mounts = [l.split() for l in open("/proc/mounts")]
    fs = max((m for m in mounts if real == m[1] or real.startswith(m[1].rstrip("/") + "/")),
             key=lambda m: len(m[1]), default=[None, None, "?"])[2]
    return view, close, f"O_TMPFILE in {d} ({fs}), MAP_SHARED, MADV_RANDOM|MADV_HUGEPAGE"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c5")
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--ndev", type=int, default=0)
    ap.add_argument("--sink", default="auto", choices=["auto", "ram", "file", "none"])
    ap.add_argument("--dir", default=None)
    ap.add_argument("--ordered", action="store_true")
    ap.add_argument("--overflow-variant", action="store_true", help="C5-ovf: 64 planted poly-W, pam500")
    ap.add_argument("--filter", action="store_true", help="apply the config's -f first (C4)")
    ap.add_argument("--samples", default=None)
    ap.add_argument("--n-samples", type=int, default=1_000_000)
    ap.add_argument("--n-rows", type=int, default=16)
    ap.add_argument("--panel", type=int, default=0)
    args = ap.parse_args()

    t0 = time.perf_counter()
    if args.overflow_variant:
        store, c = synth.make_overflow(n=args.n or 20000, planted=64)
    else:
        store, c = synth.make(args.config, n=args.n)
    gen_s = time.perf_counter() - t0
    cfg = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"), gap_open=c.get("gap_open"),
                      gap_ext=c.get("gap_ext"))
    flags = capi.SAB_FLAG_ORDERED if args.ordered else 0
    out = {"config": args.config + ("-ovf" if args.overflow_variant else ""), "sequences_in": store.num,
           "args": " ".join(synth.cli_args(**c)), "generate_s": round(gen_s, 2)}
    with MultiAligner(cfg, ndev=args.ndev, panel=args.panel, flags=flags) as ma:
        if args.filter and c.get("filter"):
            t = time.perf_counter()
            lost = ma.filter(store, float(c["filter"]))
            out["filter"] = {"threshold": c["filter"], "ms": (time.perf_counter() - t) * 1e3,
                             "lost": int(lost.sum()), "library_ms": ma.stats()["filter_ms"],
                             "matching_pairs": ma.stats()["filter_matches"],
                             "lost_checksum": int(np.flatnonzero(lost).astype(np.int64).sum())}
            store = store.subset(~lost)
        n = store.num
        pairs, cells = store.pairs(), store.cells()
        t = time.perf_counter()
        ma.upload(store)
        out["upload_s"] = round(time.perf_counter() - t, 3)
        out["devices"] = ma.ndev
        nbytes = pairs * 4
        kind = args.sink
        if kind == "auto":
            kind = "file" if 4 * n * n > 0.75 * mem_available() else "ram"
        out["mem_available_gb"] = round(mem_available() / 2**30, 1)
        view = close = None
        if kind != "none":
            t = time.perf_counter()
            view, close, desc = open_sink(kind, nbytes, args.dir)
            out["sink"] = {"kind": kind, "what": desc, "bytes": nbytes, "open_s": round(time.perf_counter() - t, 3)}
        t = time.perf_counter()
        ma.align(triangular=True, out=view, write=view is not None)
        wall = time.perf_counter() - t
        st = ma.stats()
        out.update({"sequences": n, "pairs": pairs, "cells": cells, "align_wall_s": round(wall, 3),
                    "gcups_wall": cells / wall / 1e9, "pairs_per_s_wall": pairs / wall,
                    "gcups_kernel_slowest_device": cells / ma.ndev / (st["kernel_ms"] * 1e-3) / 1e9 if st["kernel_ms"] else None,
                    "stats": st, "ordered": bool(args.ordered)})
        if view is not None:
            rng = np.random.default_rng(20261105)
            ns = min(args.n_samples, pairs)
            js = rng.integers(1, n, size=ns)
            is_ = (rng.random(ns) * js).astype(np.int64)
            idx = js * (js - 1) // 2 + is_
            t = time.perf_counter()
            vals = view[idx]
            rows = sorted(set(rng.integers(1, n, size=args.n_rows).tolist()) | {1, n - 1})
            row_vals = [np.array(view[j * (j - 1) // 2: j * (j - 1) // 2 + j]) for j in rows]
            # checksum of checksums over 64 Mi-element blocks (order-independent within a block)
            csum = 0
            blk = 1 << 26
            for o in range(0, pairs, blk):
                csum = (csum * 1000003 + int(view[o:o + blk].sum(dtype=np.int64))) % (1 << 61)
            out["readback_s"] = round(time.perf_counter() - t, 3)
            out["checksum"] = csum
            if args.samples:
                np.savez_compressed(args.samples, config=np.array(out["config"]), n=np.array(n),
                                    sequences_in=np.array(out["sequences_in"]),
                                    filtered=np.array(bool(args.filter and c.get("filter"))),
                                    i=is_.astype(np.int32), j=js.astype(np.int32), score=np.asarray(vals),
                                    rows=np.array(rows), row_scores=np.concatenate(row_vals))
                out["samples"] = {"file": args.samples, "pairs": int(ns), "rows": rows}
            del vals, row_vals
            t = time.perf_counter()
            del view
            out["sink"]["msync_close_s"] = round(close(), 3)
            out["sink"]["release_s"] = round(time.perf_counter() - t, 3)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
