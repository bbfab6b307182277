#!/usr/bin/env python3
"""bench.py -- all-vs-all GCUPS of the B200 path (BASELINE.json metric) + CPU reference arm.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c2]

One "step" = one full pass of the hot path (score every i<j pair) over the synthetic
workload.  N=1 runs BASELINE.json configs[1] (C2: 10 000 proteins, len U{250..350},
`-a ga -m blosum62 -s 10 -e 1`).  For N>1 (torchrun, one rank per GPU) the pair space is
sharded by row strips with no data-path collective; the sequence count grows with sqrt(N)
so the pairs per GPU stay fixed ("weak").

JSON keys (one line, rank 0):
  value      GCUPS, inputs resident in HBM, scores left in HBM (-W mode), device time
             (CUDA events recorded by the library on its own streams), max over ranks
  e2e        GCUPS through the C ABI with HOST buffers: sab_upload (pack + H2D) and
             sab_align into a host matrix (D2H + host scatter) inside the timed region
  roofline   dominant kernel vs the integer issue peak (DPX + IMAD rates) measured on this
             box in the same run (tools/pipe_peak.cu); HBM shown non-binding next to it
  cpu_baseline  the reference's own CPU code (oracle/_ref/seqalign) on this box's host
             cores, bounded sample of the same workload
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# algorithmic integer lane-instructions per DP cell
INSTR_PER_CELL = {"nw": 1.5, "ga": 3.0, "sw": 3.5}
# DPX (alu-pipe) instructions the kernel issues per cell
ALU_INSTR_PER_CELL = {"nw": 0.5, "ga": 1.5, "sw": 2.25}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), \
        int(os.environ.get("WORLD_SIZE", 1))


def reduce_over_ranks(ms: float, units: int):
    """max of the per-rank times, sum of the per-rank unit counts (no-op for one rank)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(ms), int(units)
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    u = torch.tensor([units], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(t.item()), int(u.item())


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"], stdout=f,
                stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        if not self.path or not os.path.exists(self.path):
            return None
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx = max(mx, float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return None
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "samples": len(sm),
                "reasons": sorted(reasons)}


def workload(config: str, n_gpus: int, n_override: int | None, method: str | None = None,
             scaling: str = "weak"):
    """The synthetic set + CLI-level configuration.  weak: the sequence count grows with sqrt(N)
    so the pairs per GPU stay fixed; strong: BASELINE.json's size at every N."""
    from sequencealigner_b200 import AlignConfig, synth
    base = synth.CONFIGS[config]
    n = n_override or (base["n"] if scaling == "strong" else int(round(base["n"] * np.sqrt(n_gpus))))
    store, c = synth.make(config, n=n)
    if method and method != c["method"]:  # informational runs: same sequences, other recurrence
        c = {k: v for k, v in c.items() if k not in ("gap_pen", "gap_open", "gap_ext")}
        c["method"] = method
        c.update({"gap_pen": 4} if method == "nw" else {"gap_open": 10, "gap_ext": 1})
    cfg = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"),
                      gap_open=c.get("gap_open"), gap_ext=c.get("gap_ext"))
    desc = {"workload": f"{config}: synthetic {'DNA' if c['alphabet'] == 'ACGT' else 'protein'} "
                        f"{n} seqs len {c['lo']}..{c['hi']}, -a {c['method']} -m {c['matrix']} "
                        + (f"-p {c['gap_pen']}" if "gap_pen" in c else
                           f"-s {c['gap_open']} -e {c['gap_ext']}")
                        + (f" -f {c['filter']}" if c.get("filter") else ""),
            "sequences": n, "pairs": store.pairs(), "cells": store.cells(),
            "seed": c["seed"], "layout": "packed-triangular int32"}
    return store, cfg, c, desc


# ---------------------------------------------------------------------------------------
# CPU reference arm / baseline
# ---------------------------------------------------------------------------------------

def cpu_reference_sample(store, c, target_s: float = 15.0):
    """Times the reference's own CPU alignment (oracle/_ref/seqalign, all host threads) on
    the first M sequences of the workload, M sized for ~target_s of CPU work.  Falls back
    to the oracle port (liboracle.so, OpenMP) when the reference binary is not there."""
    from oracle import oracle
    from sequencealigner_b200 import synth
    cores = len(os.sched_getaffinity(0)) or os.cpu_count() or 1
    cli = synth.cli_args(c["method"], c["matrix"], gap_pen=c.get("gap_pen"),
                         gap_open=c.get("gap_open"), gap_ext=c.get("gap_ext"))
    seen = {"threads": None, "binary": None}

    def run(m):
        sub = store.head(m)
        if oracle.ref_available():
            with tempfile.TemporaryDirectory() as td:
                fa = Path(td) / "sample.fasta"
                sub.to_fasta(fa)
                # threads=0 -> "-T <host cores>", OMP_* scrubbed from the child's environment
                r = oracle.run_reference(fa, cli, threads=0, write=False, bench=True)
            secs = r["align_seconds"]
            if not secs or secs <= 0:
                raise RuntimeError("could not read 'Alignment: X sec' from the reference")
            seen["threads"], seen["binary"] = r["threads"], r["binary"]
            return sub.cells(), secs, "reference"
        from sequencealigner_b200 import AlignConfig
        r = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"),
                        gap_open=c.get("gap_open"), gap_ext=c.get("gap_ext")).resolve()
        t0 = time.perf_counter()
        cells = oracle.time_columns(r, sub, 1, sub.num)
        return cells, time.perf_counter() - t0, "port"

    mean_len = float(store.lengths.mean())
    m0 = int(min(store.num, max(64, np.sqrt(2 * 4e9 / mean_len ** 2))))  # ~4e9 cells probe
    cells, secs, kind = run(m0)
    rate = cells / max(secs, 1e-3)
    m = int(min(store.num, max(m0, np.sqrt(2 * rate * target_s / mean_len ** 2))))
    if m > m0 * 1.3:
        cells, secs, kind = run(m)
    else:
        m = m0
    threads = seen["threads"] or cores
    return {"value": cells / secs / 1e9, "unit": "GCUPS", "cores": threads, "kind": kind,
            "sample": f"first {m} of {store.num} sequences ({cells:.3e} cells, {secs:.2f} s "
                      f"alignment time; the reference reports 'CPU Threads: {seen['threads']}' on "
                      f"{cores} host cores; binary oracle/_ref/{seen['binary']})",
            "_cells": cells, "_secs": secs, "_m": m}


def cpu_reference_filter_sample(store, c, m=20000):
    """The reference's own CPU filter (src/bio/filter.c, all host threads -- its OpenMP loop races
    on lost[], so only the time is used) on the first m sequences of the workload."""
    from oracle import oracle
    from sequencealigner_b200 import synth
    if not oracle.ref_available():
        return None
    sub = store.head(m)
    with tempfile.TemporaryDirectory() as td:
        fa = Path(td) / "sample.fasta"
        sub.to_fasta(fa)
        r = oracle.run_reference(fa, synth.cli_args(**c), threads=0, write=False, bench=True)
    secs = r["filter_seconds"]
    if not secs:
        return None
    return {"seconds": secs, "sequences": sub.num, "pairs_per_s": sub.pairs() / secs,
            "threads": r["threads"],
            "sample": f"first {sub.num} of {store.num} sequences, -f {c['filter']}, reference's "
                      f"'Filtering: X sec' (-B)"}


def cli_end_to_end(store, c, binary, threads=0, extra=()):
    """End-to-end pairs/s of a `seqalign` CLI build (BASELINE.json: "end-to-end pairs/sec" =
    pairs / wall time of the whole run: FASTA parse, filter, alignment, HDF5 write)."""
    import re
    import shutil
    td = tempfile.mkdtemp(prefix="sab_cli_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        fa = Path(td) / "in.fasta"
        store.to_fasta(fa)
        from sequencealigner_b200 import synth
        out = Path(td) / "out.h5"
        nthreads = threads or len(os.sched_getaffinity(0)) or 1
        cmd = [str(binary), "-i", str(fa), "-o", str(out), *synth.cli_args(**c), "-F", "-P", "-B",
               "-T", str(nthreads), *extra]
        env = {k: v for k, v in os.environ.items() if not k.startswith(("OMP_", "GOMP_", "KMP_"))}
        t0 = time.perf_counter()
        p = subprocess.run(cmd, capture_output=True, text=True, env=env)
        wall = time.perf_counter() - t0
        if p.returncode != 0:
            return {"error": (p.stdout + p.stderr)[-300:]}

        def sec(name):
            m = re.search(name + r":\s*([0-9.]+)\s*sec", p.stdout)
            return float(m.group(1)) if m else None
        dev = re.search(r"Using (\d+) B200", p.stdout)
        lost = re.search(r"Filtered out (\d+) sequences", p.stdout)  # filter.c:81
        n_after = store.num - (int(lost.group(1)) if lost else 0)
        pairs = n_after * (n_after - 1) // 2  # the pairs the run actually scored and wrote
        return {"pairs_per_s": pairs / wall, "wall_s": wall, "pairs": pairs,
                "sequences_in": store.num, "sequences_after_filter": n_after,
                "split_s": {"input": sec("Input"), "filter": sec("Filtering"),
                            "align": sec("Alignment"), "output": sec("Output")},
                "align_s": sec("Alignment"),
                "output_bytes": out.stat().st_size if out.exists() else 0,
                "devices": int(dev.group(1)) if dev else 0, "host_threads": nthreads,
                "args": " ".join(cmd[5:]),
                "what": "integration/_build/seqalign_b200 (unmodified reference host + this "
                        "library): FASTA parse + filter + alignment + HDF5 write, wall clock; "
                        "split_s = the reference's own -B timers"}
    finally:
        shutil.rmtree(td, ignore_errors=True)


def run_reference_arm(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    store, cfg, c, desc = workload(args.config, args.gpus, args.n, args.method, args.scaling)
    per_step = max(4.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_reference_sample(store, c, target_s=per_step)
    cells = secs = 0.0
    last = None
    for _ in range(args.steps):
        last = cpu_reference_sample(store, c, target_s=per_step)
        cells += last["_cells"]
        secs += last["_secs"]
    value = cells / secs / 1e9
    base = {k: v for k, v in last.items() if not k.startswith("_")}
    base["value"] = value
    out = {"impl": "reference", "metric": "all-vs-all GCUPS", "value": value, "unit": "GCUPS",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
           "scaling": args.scaling,
           "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": desc,
           "cpu_baseline": base,
           "e2e": {"value": value, "unit": "GCUPS", "h2d_bytes_per_step": 0,
                   "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)
    return 0


# ---------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------

def measure_pipe_peak(device: int):
    import ctypes as C
    from sequencealigner_b200.build import PIPE_LIB
    if not PIPE_LIB.exists():
        return None
    lib = C.CDLL(str(PIPE_LIB))
    lib.sab_pipe_peak.restype = C.c_int
    lib.sab_pipe_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    names = {0: "viaddmnmx_s16x2", 1: "vimnmx3_s16x2", 2: "viadd_16x2", 3: "iadd3", 4: "imad",
             5: "lds32_random", 6: "ga_mix_alu", 7: "ga_mix_dpx_imad", 8: "viaddmnmx_u16x2",
             9: "vimnmx3_u16x2", 10: "ga_mix_dpx_iadd3", 11: "dpx_iadd3_1to1",
             12: "lds32_conflict_free", 13: "ga_cell_dpx_iadd3_lds", 14: "ga_cell_dpx_imad_lds",
             15: "dpx_imad_1to1"}
    out = {}
    for op, name in names.items():
        g, ms = C.c_double(), C.c_double()
        if lib.sab_pipe_peak(op, device, C.byref(g), C.byref(ms)) == 0:
            out[name] = g.value / 1e3  # tera lane-instr/s
    return out


def measure_cell_ceiling(device: int, method: str, narrow: bool):
    """TCUPS of the bare 8-row register block of the kernel (tools/cell_peak.cu): what this
    instruction mix reaches on this GPU with no queue, strip boundaries or staging."""
    import ctypes as C
    from sequencealigner_b200.build import PIPE_LIB
    if not PIPE_LIB.exists():
        return None
    lib = C.CDLL(str(PIPE_LIB))
    if not hasattr(lib, "sab_cell_peak"):
        return None
    lib.sab_cell_peak.restype = C.c_int
    lib.sab_cell_peak.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    t, clk = C.c_double(), C.c_double()
    if lib.sab_cell_peak({"nw": 0, "ga": 1, "sw": 2}[method], int(narrow), device, C.byref(t), C.byref(clk)) != 0:
        return None
    return {"gcups": t.value * 1e3, "clk_per_cell_pair_per_smsp": clk.value,
            "strip_columns": 16 if narrow else 32,
            "source": "tools/cell_peak.cu measured in this run: the kernel's unrolled 8-row block alone"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from sequencealigner_b200 import Aligner

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the alignment path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world
    # --emulate-shards K: ONE GPU computes shard 0 of K of the K-GPU job (same set, same panels, no
    # neighbours): attributes per-GPU rate changes between N = 1 and N = K to the plan, not the box
    emu = args.emulate_shards if world == 1 and args.emulate_shards > 1 else 0
    nsh, my_shard = (emu, 0) if emu else (world, rank)

    store, cfg, c, desc = workload(args.config, emu or n_gpus, args.n, args.method, args.scaling)
    # `config` is kept identical to the reference arm's; what is specific to this arm goes beside it
    parallelism = f"row-strip shards x{n_gpus}, no collective"
    l2_flush = "512 MiB device memset between timed steps"
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    al = Aligner(cfg, device=local_rank, shards_hint=nsh, flags=args.flags, panel=args.panel)

    # ---- -f: the redundancy filter runs first, the alignment on the survivors ----
    filt = None
    store_in = store  # what the CLI / CPU legs are given (they filter themselves)
    if c.get("filter") and not args.no_filter:
        from sequencealigner_b200.aligner import filter_resolve
        thr = float(c["filter"])
        n_in = store.num

        def run_filter():
            if world == 1:
                return al.filter(store, thr), al.stats()["filter_matches"]
            # phase A on this rank's share of the j-tiles, the match lists exchanged over NCCL
            # (the path's one real exchange step), phase B -- sequential -- on every rank alike
            mine = torch.from_numpy(al.filter_matches(store, thr, shard=rank, nshards=world)
                                    .view(np.int64)).cuda()
            cnt = torch.tensor([mine.numel()], dtype=torch.int64, device="cuda")
            cnts = [torch.zeros_like(cnt) for _ in range(world)]
            dist.all_gather(cnts, cnt)
            cap = max(1, max(int(x.item()) for x in cnts))
            pad = torch.zeros(cap, dtype=torch.int64, device="cuda")
            pad[:mine.numel()] = mine
            parts = [torch.zeros_like(pad) for _ in range(world)]
            dist.all_gather(parts, pad)
            allp = torch.cat([p_[:int(k.item())] for p_, k in zip(parts, cnts)]).cpu().numpy()
            return filter_resolve(n_in, allp.view(np.uint64)), int(allp.size)

        run_filter()  # warm-up: device buffers, module load
        barrier()
        t_f = time.perf_counter()
        lost, n_match = run_filter()
        barrier()
        f_ms, _ = reduce_over_ranks((time.perf_counter() - t_f) * 1e3, 0)
        store = store.subset(~lost)
        filt = {"threshold": thr, "sequences_in": n_in, "sequences_after": store.num,
                "lost": int(lost.sum()), "pairs_compared": n_in * (n_in - 1) // 2,
                "matching_pairs": n_match, "ms": f_ms,
                "pairs_after": store.pairs(), "cells_after": store.cells(),
                "lost_checksum": int(np.flatnonzero(lost).astype(np.int64).sum()),
                "what": "sab_filter (1 GPU) / sab_filter_matches per rank + NCCL all_gather + "
                        "sab_filter_resolve (N GPUs): wall clock incl. packing, H2D, D2H, resolve"}
    al.upload(store)
    n = store.num
    host_matrix = np.zeros(n * (n - 1) // 2, dtype=np.int32)

    # ---- device-resident throughput (`value`) ------------------------------------------
    for _ in range(args.warmup):
        al.align(write=False, shard=my_shard, nshards=nsh)
    dev_ms, my_cells, launches = 0.0, 0, 0
    kernel_ms = 0.0
    widths = {}
    barrier()
    with ClockSampler(local_rank) as clocks:
        t_wall0 = time.perf_counter()
        for _ in range(args.steps):
            flush.zero_()
            torch.cuda.synchronize()
            al.align(write=False, shard=my_shard, nshards=nsh)
            st = al.stats()
            dev_ms += st["device_ms"]
            kernel_ms += st["kernel_ms"]
            my_cells += st["cells"]
            launches += st["launches"]
            widths = st["width_launches"]
        barrier()
        wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clk = clocks.summary()
    step_ms, step_cells = reduce_over_ranks(dev_ms / args.steps, my_cells // args.steps)
    value = step_cells / (step_ms * 1e-3) / 1e9
    kern_step_ms, _ = reduce_over_ranks(kernel_ms / args.steps, 0)
    launches_value_leg = launches

    # ---- end to end through the C ABI with host buffers (`e2e`) ------------------------
    for _ in range(min(args.warmup, 1)):
        al.upload(store)
        al.align(out=host_matrix, shard=my_shard, nshards=nsh)
    barrier()
    e2e_t0 = time.perf_counter()
    h2d = d2h = 0
    e2e_steps = []  # per step: wall ms of the two calls and the library's own device span
    for _ in range(args.steps):
        t_a = time.perf_counter()
        al.upload(store)
        t_b = time.perf_counter()
        al.align(out=host_matrix, shard=my_shard, nshards=nsh)
        t_c = time.perf_counter()
        st = al.stats()
        h2d, d2h = st["h2d_bytes"], st["d2h_bytes"]
        launches += st["launches"]
        e2e_steps.append({"upload_ms": round((t_b - t_a) * 1e3, 2), "align_ms": round((t_c - t_b) * 1e3, 2),
                          "device_ms": round(st["device_ms"], 2)})
    barrier()
    e2e_ms = (time.perf_counter() - e2e_t0) * 1e3 / args.steps
    e2e_ms, _ = reduce_over_ranks(e2e_ms, 0)
    e2e_value = step_cells / (e2e_ms * 1e-3) / 1e9
    checksum = int(host_matrix.astype(np.int64).sum())
    al.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (sab_wave_kernel<method, int16x2, K>) ------------
    # Bound: SM integer issue.  achieved = cells/s x the ALGORITHMIC
    # lane-instructions per cell of SURVEY.md section 8d; peak = DPX (alu pipe) + IMAD (fma
    # pipe) lane-instruction rates measured on this box right now by tools/pipe_peak.cu.
    peaks = measure_pipe_peak(local_rank) or {}
    method = c["method"]
    dpx_peak = max(peaks.get("viaddmnmx_u16x2", 0.0), peaks.get("vimnmx3_u16x2", 0.0)) or None
    imad_peak = peaks.get("imad") or None
    int_peak = (dpx_peak + imad_peak) if dpx_peak and imad_peak else None
    cells_per_s_gpu = step_cells / n_gpus / (kern_step_ms * 1e-3)
    achieved = cells_per_s_gpu * INSTR_PER_CELL[method] / 1e12
    alu_issued = cells_per_s_gpu * ALU_INSTR_PER_CELL[method] / 1e12
    traffic = None
    tf = ROOT / "profiles" / "traffic.json"
    if tf.exists():
        try:
            traffic = json.loads(tf.read_text()).get(f"{args.config}_bytes_per_launch")
        except (ValueError, OSError):
            traffic = None
    measured = {}
    mp = ROOT / "MEASURED_PEAKS.json"
    if mp.exists():
        try:
            measured = json.loads(mp.read_text())
        except ValueError:
            measured = {}
    hbm_peak = float(measured.get("hbm_gbs", 6650.0))
    bits = 5 if c["alphabet"] != "ACGT" else 2
    alg_bytes = float(store.pairs()) * 4 + float(store.lengths.sum()) * bits / 8 * 2
    roofline = {
        "bound": "int_alu", "kernel": f"sab_wave_kernel<{method},int16x2,K> (one launch per strip width in use)",
        "achieved": achieved, "peak": int_peak, "unit": "T lane-instr/s",
        "frac": (achieved / int_peak) if int_peak else None,
        "peak_source": "measured on this box in this run by tools/pipe_peak.cu: VIADDMNMX/VIMNMX3.U16x2 "
                       "rate (alu pipe) + IMAD rate (fma pipe), all SMs; MEASURED_PEAKS.json has no "
                       "integer figure",
        "instr_per_cell_algorithmic": INSTR_PER_CELL[method],
        "alu_pipe": {"dpx_per_cell_issued": ALU_INSTR_PER_CELL[method], "achieved": alu_issued,
                     "peak": dpx_peak, "frac": (alu_issued / dpx_peak) if dpx_peak else None,
                     "unit": "T lane-instr/s"},
        "gcups_kernel_per_gpu": cells_per_s_gpu / 1e9,
        "gcups_bound_per_gpu": (int_peak / INSTR_PER_CELL[method] * 1e3) if int_peak else None,
        "launches_per_step": launches_value_leg / args.steps,
        "pipe_peaks_T_lane_instr_s": peaks,
        "traffic": traffic,
        "hbm": {"algorithmic_bytes_per_step": alg_bytes,
                "achieved_gbs": alg_bytes / n_gpus / (kern_step_ms * 1e-3) / 1e9,
                "peak_gbs": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if measured else "fallback",
                "note": "non-binding"},
    }

    # same rule as plan_strips (sab_capi.cu): narrow strips only where they save enough padding
    ln = store.lengths.astype(np.int64)
    narrow = float(((ln + 15) // 16 * 16).sum()) * 1.055 < float(((ln + 31) // 32 * 32).sum())
    ceiling = measure_cell_ceiling(local_rank, method, narrow)
    if ceiling:
        ceiling["kernel_frac_of_ceiling"] = cells_per_s_gpu / 1e9 / ceiling["gcups"]
    roofline["code_shape_ceiling"] = ceiling

    cli = None
    dropin = ROOT / "integration" / "_build" / "seqalign_b200"
    if n_gpus == 1 and not args.no_cli and dropin.exists():
        cli = cli_end_to_end(store_in, c, dropin)

    cpu = None
    if n_gpus == 1 and not args.no_cpu:
        cpu = cpu_reference_sample(store_in, c, target_s=args.cpu_seconds)
        cpu = {k: v for k, v in cpu.items() if not k.startswith("_")}
        if filt:
            cpu["filter"] = cpu_reference_filter_sample(store_in, c)

    out = {
        "metric": "all-vs-all GCUPS", "value": value, "unit": "GCUPS", "n_gpus": n_gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "int16x2",
        "data": "synthetic", "config": desc, "parallelism": parallelism, "l2_flush": l2_flush,
        "pairs_per_s": (filt["pairs_after"] if filt else desc["pairs"]) / (step_ms * 1e-3),
        "filter": filt, "emulated_shard": (f"0 of {emu}" if emu else None),
        "wall_ms_per_step": wall_ms / args.steps,
        "e2e": {"value": e2e_value, "unit": "GCUPS", "ms_per_step": e2e_ms,
                "pairs_per_s": (filt["pairs_after"] if filt else desc["pairs"]) / (e2e_ms * 1e-3),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "checksum": checksum, "steps_rank0": e2e_steps},
        "gpu_launches": launches, "width_launches": {str(k): v for k, v in widths.items()},
        "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
        "cli_end_to_end": cli,
    }
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--n", type=int, default=None, help="override the sequence count")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: sequences x sqrt(gpus) (default, the driver's SCALE run); strong: "
                         "the config's own size at every N (BASELINE.json quotes C3 at 1/2/4/8)")
    ap.add_argument("--method", default=None, choices=["nw", "ga", "sw"],
                    help="override the config's alignment method (informational runs)")
    ap.add_argument("--emulate-shards", type=int, default=0,
                    help="one GPU computes shard 0 of K of the K-GPU job (attribution runs)")
    ap.add_argument("--no-filter", action="store_true", help="skip the -f leg of configs that have one")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-cli", action="store_true", help="skip the drop-in CLI end-to-end leg")
    ap.add_argument("--flags", type=int, default=0, help="sab_config.flags (tuning / A-B runs)")
    ap.add_argument("--panel", type=int, default=0, help="sab_config.panel (0 = auto)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    args = ap.parse_args()
    if args.steps < 1:
        raise SystemExit("--steps must be >= 1")
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
