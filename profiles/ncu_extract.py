#!/usr/bin/env python
"""Cut an `ncu -i X.ncu-rep --page raw --csv` dump down to the metrics the profile notes quote.
usage: ncu -i X.ncu-rep --page raw --csv > raw.csv ; ncu_extract.py raw.csv > profiles/<name>.csv"""
import csv, sys
KEEP = """ID
Kernel Name
gpu__time_duration.sum
launch__grid_size
launch__block_size
launch__registers_per_thread
smsp__thread_inst_executed_per_inst_executed.ratio
smsp__inst_executed.sum
smsp__issue_active.avg.pct_of_peak_sustained_active
smsp__warps_eligible.avg.per_cycle_active
sm__warps_active.avg.pct_of_peak_sustained_active
dram__bytes_read.sum
dram__bytes_write.sum
gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed
l1tex__t_sector_hit_rate.pct
lts__t_sector_hit_rate.pct
l1tex__throughput.avg.pct_of_peak_sustained_elapsed
l1tex__data_pipe_lsu_wavefronts.sum
l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum
l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_st.sum
l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum
l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum
lts__t_sectors_srcunit_tex_op_read.sum
lts__t_sectors_srcunit_tex_op_write.sum
smsp__sass_inst_executed_op_global_ld.sum
smsp__sass_inst_executed_op_global_st.sum
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio""".split("\n")
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
keep = [i for i, n in enumerate(h) if n in KEEP]
w = csv.writer(sys.stdout)
for r in rows:
    w.writerow([r[i] for i in keep])
