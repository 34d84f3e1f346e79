"""Key per-kernel numbers from `ncu -i X.ncu-rep --page raw --csv` output.  usage: python tools/ncu_raw_summary.py raw.csv"""
import csv
import sys

WANT = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__inst_executed.sum", "warp_inst"), ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1thr%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2thr%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("sm__inst_executed_pipe_fp64.sum", "fp64_inst"), ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64pipe%"),
        ("smsp__average_warp_latency_issue_stalled_long_scoreboard.pct", "st_long%"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
        ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio"),
        ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_bar"),
        ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "st_branch"),
        ("smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "st_disp"),
        ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "st_noinst")]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
cols = [(hdr.index(n), s) for n, s in WANT if n in hdr]
for r in rows[2:]:
    out = []
    for i, s in cols:
        v = r[i]
        if s == "kernel":
            v = v.replace("<unnamed>::", "").split("(")[0][:34]
        else:
            try:
                v = "%.4g" % float(v.replace(",", ""))
            except ValueError:
                pass
            v = v + (units[i] if s in ("time", "dram_rd", "dram_wr") else "")
        out.append("%s=%s" % (s, v))
    print("  ".join(out))
