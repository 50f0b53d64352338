#!/usr/bin/env python3
"""Condense an `ncu --set full` report into the few numbers DESIGN.md / profiles/README.md quote.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [units-per-launch] > profiles/x_ncu.csv

Reads the report with `ncu -i ... --page raw --csv` and `--page source --csv` (instruction mix and stall samples per
opcode).  `units` (optional) = work units of the captured launch, to print instructions per unit."""
import collections
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else None
    raw = page(rep, "raw")
    hdr, unit, val = raw[0], raw[1], raw[2]
    w = csv.writer(sys.stdout)
    w.writerow(["metric", "unit", "value"])
    name_i = hdr.index("Kernel Name") if "Kernel Name" in hdr else None
    if name_i is not None:
        w.writerow(["kernel", "", val[name_i]])
    for h, u, v in zip(hdr, unit, val):
        if h in KEEP or h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            try:
                if h.startswith("smsp__average") and float(v) < 0.05:
                    continue
            except ValueError:
                pass
            w.writerow([h, u, v])
    src = page(rep, "source")
    if len(src) > 2:
        h2 = src[1]
        i_s, i_n, i_x = h2.index("Source"), h2.index("# Samples"), h2.index("Instructions Executed")
        samp, execd = collections.Counter(), collections.Counter()
        for r in src[2:]:
            if len(r) <= i_x:
                continue
            ins = r[i_s].strip().split()
            if not ins:
                continue
            op = (ins[1] if ins[0].startswith("@") else ins[0]).split(".")[0].rstrip(";")
            samp[op] += int(r[i_n] or 0); execd[op] += int(r[i_x] or 0)
        ts, tx = sum(samp.values()) or 1, sum(execd.values()) or 1
        w.writerow(["instructions_executed_total", "warp inst", tx])
        if units:
            w.writerow(["instructions_per_unit", "warp inst", round(tx / units, 1)])
        for op, n in execd.most_common(14):
            w.writerow([f"opcode {op}", "% of executed / % of stall samples" + (" / per unit" if units else ""),
                        f"{100 * n / tx:.1f} / {100 * samp[op] / ts:.1f}" + (f" / {n / units:.1f}" if units else "")])


if __name__ == "__main__":
    main()
