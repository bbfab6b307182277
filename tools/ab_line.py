#!/usr/bin/env python3
"""One summary line of a bench.py JSON line read from stdin (tools/ab2.sh)."""
import json
import sys

label = " ".join(sys.argv[1:])
lines = [l for l in sys.stdin.read().strip().splitlines() if l.startswith("{")]
if not lines:
    print(label, "| no JSON line")
    sys.exit(0)
d = json.loads(lines[-1])
r = d.get("roofline") or {}
print(label, "|", d["config"]["workload"][:48],
      "value=%.0f e2e=%.0f frac=%.3f kernel/gpu=%.0f widths=%s checksum=%s" %
      (d["value"], d["e2e"]["value"], r.get("frac") or 0, r.get("gcups_kernel_per_gpu") or 0,
       d.get("width_launches"), d["e2e"].get("checksum")))
