#!/usr/bin/env python3
"""Prints selected metrics from `ncu --page raw --csv` output (one column per profiled launch)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
pats = sys.argv[2:] or ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sectors_op_read.sum', 'lts__t_bytes.sum',
  'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active',
  'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
  'smsp__thread_inst_executed.sum', 'launch__registers_per_thread', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
  'l1tex__throughput.avg.pct_of_peak_sustained_active', 'smsp__sass_average_branch_targets_threads_uniform.pct', 'sm__cycles_elapsed.avg',
  'smsp__cycles_active.avg', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
  'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_lsu.sum',
  'smsp__warps_eligible.avg.per_cycle_active', 'smsp__inst_executed_op_local_ld.sum', 'smsp__inst_executed_op_local_st.sum',
  'smsp__average_warp_latency_issue_stalled', 'smsp__average_warps_issue_stalled']
for i, h in enumerate(hdr):
    if any(p in h for p in pats):
        print(h, '[%s]' % units[i], ' | '.join(r[i] for r in rows[2:]))
