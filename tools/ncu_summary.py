#!/usr/bin/env python
"""Selected metrics of every kernel of an `ncu --page raw --csv` export, one block per launch
(the text summaries under profiles/).  usage: ncu_summary.py RAW.csv > summary.txt"""
import csv
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
    "lts__t_sector_hit_rate.pct", "smsp__inst_executed_op_shared_atom.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    ti = col["gpu__time_duration.sum"]
    to_us = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(units[ti], 1.0)
    num = lambda x: float(x.replace(",", "")) if x not in ("", "n/a") else 0.0     # noqa: E731
    total = sum(num(r[ti]) * to_us for r in rows[2:])
    for r in rows[2:]:
        name = r[col["Kernel Name"]].replace("pmb::", "")
        print("== %s   (%.1f %% of the captured kernels' time)" % (name, 100.0 * num(r[ti]) * to_us / total))
        for m in METRICS:
            if m in col:
                print("   %-88s %s %s" % (m, r[col[m]], units[col[m]]))
        print()


if __name__ == "__main__":
    main()
