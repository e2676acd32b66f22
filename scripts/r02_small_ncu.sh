#!/bin/bash
O=gpurun_out; mkdir -p $O; T=${1:-r02j}
timeout 120 python scripts/gather_small.py > $O/${T}_gs.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gather_small_kernel -s 2 -c 1 -f -o $O/${T}_gather_smallk \
   python scripts/gather_small.py > $O/${T}_gs_ncu.log 2>&1
cat $O/${T}_gs.log
