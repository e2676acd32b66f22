#!/bin/bash
# launch list of a whole-program frame (cornell 256^2 aa 1) -> gpurun_out/$1_launches_frame_cornell256aa1.csv
O=gpurun_out; mkdir -p $O; T=${1:-r02b}
REPS=1 timeout 300 python scripts/render_perf.py cornell 256 1 > $O/${T}_f_small.log 2>&1 && \
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/${T}_launches_frame_cornell256aa1.csv \
    python scripts/render_perf.py cornell 256 1 > $O/${T}_ncu_f.log 2>&1
cat $O/${T}_f_small.log
