#!/usr/bin/env python
"""Warp-instructions executed per CUDA source line for one kernel:
   python tools/ncu_lines.py <ncu --page source --csv dump> <cubin> <kernel-substring> [top]
Joins the per-SASS-instruction counts of the ncu dump with `nvdisasm -g -c` line markers (same order)."""
import collections
import csv
import re
import subprocess
import sys


def main(dump, cubin, kernel, top=40):
    rows = list(csv.reader(open(dump)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    h = rows[hi]
    ei = h.index("Instructions Executed")
    counts = [int(r[ei]) for r in rows[hi + 1:] if len(r) > ei and r[ei].isdigit()]
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    # isolate the kernel's section
    sec = re.split(r"\n\s*\.section\s+\.text\.", txt)
    body = [s for s in sec if kernel in s.split("\n", 1)[0]]
    assert body, "kernel not found"
    body = body[0] if len(body) == 1 else [b for b in body if "Lb1" in b.split("\n", 1)[0]][0]
    line, seq = ("?", 0), []
    for ln in body.split("\n"):
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            line = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s", ln):
            seq.append(line)
    print(f"sass instructions: dump {len(counts)}, disasm {len(seq)}")
    n = min(len(counts), len(seq))
    per = collections.Counter()
    for i in range(n):
        per[seq[i]] += counts[i]
    tot = sum(per.values())
    for (f, l), c in per.most_common(int(top)):
        print(f"{f}:{l:<5d} {c:12d} {c / tot:6.3f}")


if __name__ == "__main__":
    main(*sys.argv[1:])
