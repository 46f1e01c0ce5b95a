#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump of ONE kernel: executed instructions by opcode and
warp-stall samples by reason.  python tools/ncu_stalls.py file.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr, data = rows[hi], rows[hi + 1:]
nxt = [i for i, r in enumerate(data) if r and r[0] in ("Kernel Name", "Address")]
if nxt:
    data = data[:nxt[0]]          # several launches in one dump: summarise the first
col = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
ops, stalls, tot_inst, tot_thr = collections.Counter(), collections.Counter(), 0, 0
for r in data:
    if len(r) < len(hdr):
        continue
    src = r[col["Source"]].strip().split()
    op = next((t for t in src if not t.startswith("@")), "?").split(".")[0]
    n = int(r[col["Instructions Executed"]] or 0)
    ops[op] += n
    tot_inst += n
    tot_thr += int(r[col["Thread Instructions Executed"]] or 0)
    for s in stall_cols:
        stalls[s] += int(r[col[s]] or 0)
print(f"warp instructions executed: {tot_inst}   thread instructions: {tot_thr}")
print("top opcodes:", ", ".join(f"{o} {100.0 * n / tot_inst:.1f}%" for o, n in ops.most_common(14)))
ts = sum(stalls.values())
print("stall samples:", ", ".join(f"{s[6:]} {100.0 * n / ts:.1f}%" for s, n in stalls.most_common(8)))
