#!/usr/bin/env python
"""Key metrics + SASS opcode histogram of one `ncu --set full` capture, as markdown:
   python tools/ncu_full_summary.py gpurun_out/full_X.ncu-rep > profiles/rNN_X_full.md"""
import csv
import io
import subprocess
import sys

sys.path.insert(0, __file__.rsplit("/", 1)[0])
import ncu_opcodes  # noqa: E402

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, v = rows[0], rows[1], rows[2]
    print(f"# `{v[h.index('Kernel Name')]}`\n")
    print(f"source: `{rep}` (`ncu --set full --clock-control none --import-source on`, one launch)\n")
    print("| metric | value | unit |\n|---|---:|---|")
    for w in WANT:
        if w in h:
            print(f"| {w} | {v[h.index(w)]} | {u[h.index(w)]} |")
    print("\nwarp stall reasons (issue-stalled warps per issue-active cycle, > 0.15):\n")
    for i, n in enumerate(h):
        if "issue_stalled" in n and n.endswith("per_issue_active.ratio"):
            try:
                if float(v[i]) > 0.15:
                    print(f"* {n.split('issue_stalled_')[1].split('_per_issue')[0]}: {float(v[i]):.2f}")
            except ValueError:
                pass
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    tmp = "/tmp/_ncu_src.csv"
    open(tmp, "w").write(src)
    print("\nSASS opcode histogram (warp-instructions executed):\n\n```")
    ncu_opcodes.main(tmp, 22)
    print("```")


if __name__ == "__main__":
    main(sys.argv[1])
