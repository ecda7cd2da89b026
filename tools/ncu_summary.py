#!/usr/bin/env python
"""tools/ncu_summary.py REPORT.ncu-rep [title] -- the metrics this repository quotes, per profiled launch, as text for profiles/.
Reads the report with `ncu -i ... --page raw --csv` (works without a GPU)."""
import csv
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
]


def main():
    report = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else report
    text = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(text.splitlines()))
    header, units = rows[0], rows[1]
    col = {name: i for i, name in enumerate(header)}
    print("# %s" % title)
    for row in rows[2:]:
        print("## %s" % row[col["Kernel Name"]])
        for name in KEEP:
            if name in col:
                print("%-80s %-16s %s" % (name, units[col[name]], row[col[name]]))
        stalls = []
        for name, i in col.items():
            if "issue_stalled" in name and name.endswith("_per_issue_active.ratio") and "not_issued" not in name:
                try:
                    stalls.append((float(row[i]), name.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        print("stall cycles per issued instruction (warp states sampled per issue):")
        for value, name in sorted(stalls, reverse=True)[:12]:
            print("    %-24s %.3f" % (name, value))
        print()


if __name__ == "__main__":
    main()
