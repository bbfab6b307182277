#!/usr/bin/env python3
"""Per-kernel SASS opcode histogram of a .so/.cubin (cuobjdump -sass); optional regex filter.
usage: tools/sass_hist.py <file> [kernel-regex] [--loop]  (--loop: only the hottest loop body,
i.e. the instructions between the last backward branch target and the branch)"""
import collections
import re
import subprocess
import sys


def main():
    path = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else None
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    cur, hist = None, {}
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            hist[cur][m.group(1)] += 1
    for k, h in hist.items():
        if pat and not pat.search(k):
            continue
        print(f"== {k}  ({sum(h.values())} instr)")
        for op, n in h.most_common(14):
            print(f"   {n:6d} {op}")


if __name__ == "__main__":
    main()
