#!/bin/sh
# usage: exp/ncu_sum.sh file.ncu-rep  -> key metrics per launch
ncu -i "$1" --page raw --csv 2>/dev/null | python3 -c '
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr=rows[0]; units=rows[1]; data=rows[2:]
want=["Kernel Name","gpu__time_duration.sum","launch__registers_per_thread","sm__warps_active.avg.pct_of_peak_sustained_active","smsp__thread_inst_executed_per_inst_executed.ratio","smsp__issue_active.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","l1tex__t_sector_hit_rate.pct","lts__t_sector_hit_rate.pct","dram__bytes_read.sum","dram__bytes_write.sum","l1tex__data_pipe_lsu_wavefronts.sum","sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active","l1tex__throughput.avg.pct_of_peak_sustained_active","lts__throughput.avg.pct_of_peak_sustained_elapsed","smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio","smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio","smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct","smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct","smsp__warp_issue_stalled_wait_per_warp_active.pct","smsp__warp_issue_stalled_not_selected_per_warp_active.pct","smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct","smsp__warp_issue_stalled_no_instruction_per_warp_active.pct","smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct","smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct"]
for w in want:
    if w in hdr:
        i=hdr.index(w)
        print(w[:70].ljust(70), units[i][:8].ljust(8), " | ".join((r[i][:14] if w!="Kernel Name" else r[i][:14]) for r in data))
'
