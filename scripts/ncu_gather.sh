#!/bin/bash
# One ncu --set full capture of the gather kernel on the small bench step (512x512 aa 1 => 1048576 gathers, k = 225,
# 1e7-photon map), after the same command exited 0 without ncu.  Output: gpurun_out/<tag>_gather.ncu-rep
TAG=${1:-gather}
CMD="python bench.py --steps 1 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --no-frame"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gather_kernel -s 1 -c 1 -f -o gpurun_out/${TAG}_gather $CMD > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_plain.log | cut -c1-400
