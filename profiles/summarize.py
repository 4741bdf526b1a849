#!/usr/bin/env python
"""Summarise an Nsight Compute report (read here, on the CPU box) into the few numbers DESIGN.md cites.

    python profiles/summarize.py gpurun_out/prof.ncu-rep [kernel-regex] > profiles/<name>.md
"""
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
]


def main():
    rep = sys.argv[1]
    pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kcol = hdr.index("Kernel Name")
    print(f"# ncu summary of `{rep}`\n")
    for r in data:
        if pat and not pat.search(r[kcol]):
            continue
        print(f"## {r[kcol][:100]}\n")
        print("| metric | unit | value |\n|---|---|---|")
        for i, h in enumerate(hdr):
            stall = "issue_stalled" in h and h.endswith("per_issue_active.ratio")
            if h in KEYS or stall:
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if stall and v < 0.1:
                    continue
                print(f"| {h} | {units[i]} | {v:.6g} |")
        print()


if __name__ == "__main__":
    main()
