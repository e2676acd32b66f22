#!/bin/bash
# ncu --set full of one launch of a kernel (regex $1, skip $2) inside a cornell 256^2 aa 1 frame -> gpurun_out/$3.ncu-rep
mkdir -p gpurun_out
REPS=1 timeout 120 python scripts/render_perf.py cornell 256 1 > gpurun_out/$3_plain.log 2>&1 && \
REPS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 1 -f -o gpurun_out/$3 \
   python scripts/render_perf.py cornell 256 1 > gpurun_out/$3_ncu.log 2>&1
cat gpurun_out/$3_plain.log
