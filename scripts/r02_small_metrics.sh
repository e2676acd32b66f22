#!/bin/bash
O=gpurun_out; mkdir -p $O; T=${1:-r02t}
python scripts/gather_small.py 2900 > $O/${T}_gs.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,sm__cycles_active.min,sm__cycles_active.max --clock-control none -k regex:small -s 4 -c 2 --csv --log-file $O/${T}_small_kernels.csv python scripts/gather_small.py 2900 > $O/${T}_ncu.log 2>&1
GI_GATHER_DEBUG=1 python scripts/gather_small.py 2900 2>&1 | grep "gather_small\]" | tail -1
