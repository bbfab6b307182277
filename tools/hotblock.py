#!/usr/bin/env python3
"""Opcode histogram of the largest straight-line blocks of each kernel in a cubin/.so
(cuobjdump -sass): the unrolled DP row-group code.  usage: tools/hotblock.py <file> [kernel-regex] [nblocks]"""
import collections
import re
import subprocess
import sys


def main():
    path = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
    nb = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    kernels, cur = {}, None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            kernels[cur].append((int(m.group(1), 16), m.group(2).strip()))
    for name, ins in kernels.items():
        if pat and not pat.search(name):
            continue
        targets = set()
        for a, s in ins:
            m = re.search(r"\b(?:BRA|BSSY\S*|CALL\S*)\b.*?(0x[0-9a-f]+)", s)
            if m and s.split()[0 if not s.startswith("@") else 1].startswith("BRA"):
                targets.add(int(m.group(1), 16))
        blocks, curb = [], []
        for a, s in ins:
            if a in targets and curb:
                blocks.append(curb)
                curb = []
            curb.append(s)
            op = re.sub(r"^@!?U?P\d+\s+", "", s).split()[0]
            if op.startswith(("BRA", "EXIT", "BAR", "RET")):
                blocks.append(curb)
                curb = []
        if curb:
            blocks.append(curb)
        blocks.sort(key=len, reverse=True)
        print(f"== {name}: {len(ins)} instr, {len(blocks)} blocks")
        for b in blocks[:nb]:
            h = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", s).split()[0] for s in b)
            dpx = sum(v for k, v in h.items() if k.startswith(("VIADDMNMX", "VIMNMX3")))
            print(f"  block of {len(b)} instr, DPX {dpx}: " +
                  ", ".join(f"{k} {v}" for k, v in h.most_common(9)))


if __name__ == "__main__":
    main()
