"""Condense `ncu -i X.ncu-rep --page raw --csv` output into the handful of counters DESIGN.md argues with.

  ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/x.csv
  python scripts/ncu_summary.py /tmp/x.csv > profiles/rNN_X_summary.csv
"""
import csv
import sys

WANT = [
    "Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
]

rows = list(csv.reader(open(sys.argv[1])))
header, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(header)}
out = csv.writer(sys.stdout)
out.writerow(["metric", "unit"] + ["launch %d" % i for i in range(len(rows) - 2)])
for name in WANT:
    if name in col:
        out.writerow([name, units[col[name]]] + [r[col[name]] for r in rows[2:]])
