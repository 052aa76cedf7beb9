"""Summary CSV (metric,unit,value) of one kernel launch in an ncu report — the files under profiles/*_ncu_full_summary.csv
(development helper).

  python tools/ncu_summary.py REPORT.ncu-rep KERNEL_REGEX [launch_skip] > profiles/rNN_ncu_full_summary.csv
"""
import csv
import io
import subprocess
import sys

METRICS = """gpu__time_duration.sum dram__bytes_read.sum dram__bytes_write.sum launch__registers_per_thread launch__block_size
launch__grid_size launch__shared_mem_per_block_dynamic launch__occupancy_limit_shared_mem
sm__warps_active.avg.pct_of_peak_sustained_active smsp__thread_inst_executed_per_inst_executed.ratio
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active
sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio sm__icc_request_hit_rate.pct
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum l1tex__t_sector_hit_rate.pct lts__t_sector_hit_rate.pct
dram__throughput.avg.pct_of_peak_sustained_elapsed""".split()


def main():
    rep, kre = sys.argv[1], sys.argv[2]
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", f"regex:{kre}", "--launch-skip", skip,
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, val = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(hdr)}
    w = csv.writer(sys.stdout, lineterminator="\n")
    w.writerow(["metric", "unit", "value"])
    w.writerow(["Kernel Name", "", val[col["Kernel Name"]]])
    for m in METRICS:
        if m in col:
            w.writerow([m, units[col[m]], val[col[m]]])


if __name__ == "__main__":
    main()
