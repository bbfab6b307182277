#!/usr/bin/env python3
"""Summarises an .ncu-rep (raw page) into the handful of metrics DESIGN.md / profiles/ cite.
usage: tools/ncu_summary.py <report.ncu-rep> [--md]"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__average_warp_latency_per_inst_issued.ratio",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        print(f"## {d.get('Kernel Name')}  (launch id {d.get('ID')})")
        for k in KEYS:
            if k in d:
                print(f"- {k} = {d[k]} {u[k]}")
        stalls = [(float(v), k) for k, v in d.items()
                  if k.startswith("smsp__average_warps_issue_stalled_") and
                  k.endswith("_per_issue_active.ratio") and v not in ("", "n/a")]
        print("- stall reasons (warps per issue-active cycle): " + ", ".join(
            f"{k.split('stalled_')[1].split('_per_')[0]} {v:.3f}" for v, k in sorted(stalls, reverse=True)[:8]))


if __name__ == "__main__":
    main()
