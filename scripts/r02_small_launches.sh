#!/bin/bash
O=gpurun_out; mkdir -p $O; T=${1:-r02m}
timeout 120 python scripts/gather_small.py > $O/${T}_gs.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio --clock-control none -k regex:small -c 8 --csv --log-file $O/${T}_small_kernels.csv python scripts/gather_small.py > $O/${T}_gs_ncu.log 2>&1
cat $O/${T}_gs.log
