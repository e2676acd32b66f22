#!/bin/bash
O=gpurun_out; mkdir -p $O; T=${1:-r02g}
timeout 300 python scripts/kd_build_only.py > $O/${T}_kd.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/${T}_launches_kd_build.csv python scripts/kd_build_only.py > $O/${T}_kd_ncu.log 2>&1
cat $O/${T}_kd.log
