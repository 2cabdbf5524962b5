"""Print selected metrics from an .ncu-rep (first kernel) -- usage: python profiles/ncu_metrics.py rep [substr ...]"""
import csv, subprocess, sys
rep = sys.argv[1]
want = sys.argv[2:] or ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct",
        "sm__pipe_fma_cycles_active.avg.pct", "sm__pipe_tensor", "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct",
        "launch__registers_per_thread", "launch__occupancy_limit", "sm__inst_executed_pipe_lsu.avg.pct",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct", "smsp__average_warp", "smsp__warp_issue_stalled",
        "sm__inst_executed_pipe_xu", "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct", "launch__shared_mem_per_block",
        "launch__grid_size", "sm__cycles_elapsed.max"]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
for i, h in enumerate(hdr):
    if any(w in h for w in want):
        print(f"{h:90s} {vals[i]:>16s} {units[i]}")
