#!/usr/bin/env python
"""Summarise an `ncu --csv` launch list (gpu__time_duration.sum [+ dram__bytes_*]) per kernel:
   python tools/ncu_summary.py gpurun_out/launches.csv > profiles/rNN_<what>_launches.md"""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    ki, mi, vi, ui = (hdr.index(c) for c in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
    per = collections.OrderedDict()
    for r in data:
        if len(r) > vi:
            per.setdefault((r[0], r[ki]), {})[r[mi]] = (float(r[vi].replace(",", "")), r[ui])
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for (_, k), m in per.items():
        name = k.split("(")[0].replace("void ", "").replace("ash::", "")
        val = lambda key: m[key][0] * scale.get(m[key][1], 1) if key in m else 0.0
        a = agg[name]
        a[0] += 1
        a[1] += val("gpu__time_duration.sum")
        a[2] += val("dram__bytes_read.sum")
        a[3] += val("dram__bytes_write.sum")
    tot = sum(a[1] for a in agg.values())
    print(f"source: {path} — per-launch times under ncu are cold-cache and serialised: compare SHARES\n")
    print("| kernel | launches | avg us | share of GPU time | DRAM read MB/launch | DRAM write MB/launch |")
    print("|---|---:|---:|---:|---:|---:|")
    for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{n}` | {a[0]} | {a[1] / a[0]:.1f} | {a[1] / tot:.3f} | {a[2] / a[0] / 1e6:.2f} | {a[3] / a[0] / 1e6:.2f} |")


if __name__ == "__main__":
    main(sys.argv[1])
