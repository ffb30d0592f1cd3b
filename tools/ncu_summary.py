#!/usr/bin/env python
"""Key metrics of an `ncu --set full` report as text (run where ncu is installed; no GPU needed).
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [> profiles/rNN_ncu_<what>.txt]"""
import csv
import subprocess
import sys

KEYS = """launch__grid_size launch__block_size launch__registers_per_thread launch__occupancy_limit_registers
launch__occupancy_limit_shared_mem launch__waves_per_multiprocessor gpu__time_duration.sum sm__cycles_elapsed.max
sm__cycles_active.avg dram__bytes_read.sum dram__bytes_write.sum gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed
lts__t_sector_hit_rate.pct lts__throughput.avg.pct_of_peak_sustained_elapsed l1tex__throughput.avg.pct_of_peak_sustained_elapsed
sm__throughput.avg.pct_of_peak_sustained_elapsed sm__warps_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum
smsp__issue_active.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_alu.sum sm__inst_executed_pipe_fma.sum sm__inst_executed.sum
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed
l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_membar_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio""".split()

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
print("# %s  (ncu --set full --clock-control none; %d launch(es) captured)" % (rep.split("/")[-1], len(data)))
ki = hdr.index("Kernel Name")
print("kernel: " + data[0][ki][:100])
print("kernels: " + " | ".join(r[ki].split("(")[0].replace("<unnamed>::", "").replace("void ", "")[:60] for r in data))
for k in KEYS:
    if k in hdr:
        i = hdr.index(k)
        print("%-86s %-16s %s" % (k, units[i], "  ".join(r[i] for r in data)))
