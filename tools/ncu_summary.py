"""Condense `ncu -i X.ncu-rep --page raw --csv` (stdin or file) into a short metric,unit,value table.

    ncu -i gpurun_out/r1_v10_rung_c2.ncu-rep --page raw --csv | python tools/ncu_summary.py > profiles/NAME.csv
"""
import csv
import sys

WANT = [
    "Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def main():
    f = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
    rows = [r for r in csv.reader(f) if len(r) > 10]
    hdr, units = rows[0], rows[1]
    out = csv.writer(sys.stdout)
    out.writerow(["launch", "metric", "unit", "value"])
    for n, vals in enumerate(rows[2:]):
        for i, h in enumerate(hdr):
            if h in WANT:
                out.writerow([n, h, units[i], vals[i]])


if __name__ == "__main__":
    main()
