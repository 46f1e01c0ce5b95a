#!/usr/bin/env python
"""Opcode histogram of one kernel from `ncu -i X.ncu-rep --page source --csv`:
   python tools/ncu_opcodes.py dump.csv [top]   -> warp-instructions executed per SASS opcode."""
import collections
import csv
import sys


def main(path, top=30):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    h = rows[hi]
    si, ei, st = h.index("Source"), h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
    per, stall = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) <= ei or not r[ei].isdigit():
            continue
        toks = r[si].split()
        op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
        op = ".".join(op.split(".")[:2]) if op.startswith(("F2F", "MUFU", "LDS", "STS", "LDG", "STG", "F2I", "I2F")) else op.split(".")[0]
        per[op] += int(r[ei])
        stall[op] += int(r[st]) if r[st].isdigit() else 0
    tot, stot = sum(per.values()), max(sum(stall.values()), 1)
    print(f"total warp-instructions {tot}")
    for op, n in per.most_common(int(top)):
        print(f"{op:18s} {n:12d} {n / tot:6.3f}   stall-samples {stall[op] / stot:6.3f}")


if __name__ == "__main__":
    main(*sys.argv[1:])
