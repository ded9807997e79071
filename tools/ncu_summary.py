"""Per-kernel summary of an ncu report (developer tool): time, DRAM bytes, issue, occupancy.
    ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram read"),
        ("dram__bytes_write.sum", "dram write"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("launch__registers_per_thread", "registers"), ("launch__grid_size", "grid"),
        ("launch__block_size", "block"), ("launch__shared_mem_per_block_dynamic", "dyn smem"),
        ("smsp__inst_executed.sum", "warp instructions"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
        ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long scoreboard"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short scoreboard"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math throttle"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg throttle"),
        ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "stall mio throttle"),
        ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall membar"),
        ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not selected")]
for r in rows[2:]:
    print("-" * 100)
    for key, label in want:
        if key in idx and r[idx[key]] != "":
            print(f"  {label:24s} {r[idx[key]][:110]} {units[idx[key]]}")
