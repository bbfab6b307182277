#!/usr/bin/env python3
"""Re-scores the samples tools/full_run.py saved on the GPU box with the CPU oracle (TEST
INFRASTRUCTURE: runs in the build container, off the GPU clock) and reports mismatches.

    python tools/verify_samples.py gpurun_out/c5.npz [--max-pairs N]
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import oracle  # noqa: E402
from sequencealigner_b200 import AlignConfig, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("npz")
    ap.add_argument("--max-pairs", type=int, default=0)
    args = ap.parse_args()
    z = np.load(args.npz)
    config = str(z["config"])
    n_in = int(z["sequences_in"])
    if config.endswith("-ovf"):
        store, c = synth.make_overflow(n=n_in, planted=64)
    else:
        store, c = synth.make(config, n=n_in)
    out = {"file": args.npz, "config": config, "sequences_in": n_in}
    if bool(z["filtered"]):
        t = time.perf_counter()
        lost = oracle.filter_lost(store, float(c["filter"]))
        store = store.subset(~lost)
        out["oracle_filter_s"] = round(time.perf_counter() - t, 1)
        out["oracle_lost"] = int(lost.sum())
    assert store.num == int(z["n"]), (store.num, int(z["n"]))
    r = AlignConfig(c["method"], c["matrix"], gap_pen=c.get("gap_pen"), gap_open=c.get("gap_open"),
                    gap_ext=c.get("gap_ext")).resolve()
    i, j, score = z["i"], z["j"], z["score"]
    if args.max_pairs and i.size > args.max_pairs:
        i, j, score = i[:args.max_pairs], j[:args.max_pairs], score[:args.max_pairs]
    t = time.perf_counter()
    want = oracle.pairs(r, store, i, j)
    out["pairs_checked"] = int(i.size)
    out["pair_mismatches"] = int((want != score).sum())
    out["max_score"] = int(score.max())
    rows, flat = z["rows"], z["row_scores"]
    pos = bad_rows = cells = 0
    for jrow in rows.tolist():
        got = flat[pos:pos + jrow]
        pos += jrow
        ii = np.arange(jrow, dtype=np.int32)
        w = oracle.pairs(r, store, ii, np.full(jrow, jrow, dtype=np.int32))
        bad_rows += int((w != got).sum())
        cells += jrow
    out["rows_checked"] = rows.tolist()
    out["row_cells_checked"] = cells
    out["row_mismatches"] = bad_rows
    out["oracle_s"] = round(time.perf_counter() - t, 1)
    print(json.dumps(out))
    return 0 if out["pair_mismatches"] == 0 and bad_rows == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
