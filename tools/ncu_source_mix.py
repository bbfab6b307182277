#!/usr/bin/env python3
"""Executed-instruction mix and stall samples per opcode from an .ncu-rep source page.
usage: tools/ncu_source_mix.py <report.ncu-rep> [top-N]"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep = sys.argv[1]
    topn = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    # a report may hold several kernels: split at "Kernel Name" rows
    k = 0
    while k < len(rows):
        if rows[k] and rows[k][0] == "Kernel Name":
            name = rows[k][1]
            hdr = rows[k + 1]
            k += 2
            body = []
            while k < len(rows) and not (rows[k] and rows[k][0] == "Kernel Name"):
                body.append(rows[k])
                k += 1
            report(name, hdr, body, topn)
        else:
            k += 1


def report(name, hdr, body, topn):
    ix = {h: i for i, h in enumerate(hdr)}
    ex, samp = collections.Counter(), collections.Counter()
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    stalls = collections.Counter()
    for r in body:
        if len(r) < len(hdr):
            continue
        m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_x]+)*)", r[ix["Source"]])
        if not m:
            continue
        op = m.group(1).split(".")[0]
        if op in ("VIADDMNMX", "VIMNMX3", "VIMNMX", "VIADD"):
            op = m.group(1).split(".")[0] + "." + ".".join(m.group(1).split(".")[1:2])
        ex[op] += int(r[ix["Instructions Executed"]] or 0)
        samp[op] += int(r[ix["# Samples"]] or 0)
        for c in stall_cols:
            stalls[c] += int(r[ix[c]] or 0)
    tot, tots = sum(ex.values()), sum(samp.values())
    print(f"## {name}\n- warp instructions executed: {tot}")
    for op, n in ex.most_common(topn):
        print(f"  - {op:<18} {n:>14}  {100.0 * n / tot:5.1f} %   samples {100.0 * samp[op] / max(tots, 1):5.1f} %")
    ts = sum(stalls.values())
    print("- stall samples: " + ", ".join(f"{c[6:]} {100.0 * v / ts:.1f}%" for c, v in stalls.most_common(8)))


if __name__ == "__main__":
    main()
