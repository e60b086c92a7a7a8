#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the few counters DESIGN.md argues from.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_<name>.txt"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_per_inst_executed.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__sass_average_branch_targets_threads_uniform.pct",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"# {rep}: ncu --set full --clock-control none (values per launch; cold-cache, serialised)")
    for r in rows[2:]:
        print(f"kernel: {r[hdr.index('Kernel Name')]}")
        for k in KEYS:
            if k in hdr:
                print(f"  {k:95s} {r[hdr.index(k)]:>18s} {units[hdr.index(k)]}")
        try:  # FP32 flops executed = (FADD + FMUL + 2 FFMA thread instructions per elapsed cycle) x elapsed cycles
            g = lambda k: float(r[hdr.index(k)])
            cyc = g("smsp__cycles_elapsed.max")
            fl = (g("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") +
                  g("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed") +
                  2 * g("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed")) * cyc
            print(f"  {'derived: fp32 flops executed (fadd + fmul + 2 ffma)':95s} {fl:18.0f} flop  ({cyc:.0f} cycles elapsed)")
        except (ValueError, KeyError):
            pass
        print()


if __name__ == "__main__":
    main()
