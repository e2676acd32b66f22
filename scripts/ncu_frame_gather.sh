#!/bin/bash
# ncu --set full of one gather launch inside a whole-program frame (cornell 256^2 aa 1: the launches are indirect-light
# gathers on the 2176-photon global map, k = 50).  Output: gpurun_out/<tag>_framegather.ncu-rep
TAG=${1:-frame}
mkdir -p gpurun_out
REPS=1 timeout 120 python scripts/render_perf.py cornell 256 1 > gpurun_out/${TAG}_plain.log 2>&1 && \
REPS=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:gather_kernel -s ${SKIP:-6} -c 1 -f -o gpurun_out/${TAG}_framegather \
   python scripts/render_perf.py cornell 256 1 > gpurun_out/${TAG}_ncu.log 2>&1
cat gpurun_out/${TAG}_plain.log
