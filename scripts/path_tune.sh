#!/bin/bash
mkdir -p gpurun_out
for r in 3 4 6; do
  GI_PATH_PERSISTENT=0 GI_PATH_OCC=$r REPS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/pt_$r.csv python scripts/render_perf.py ${1:-cornell} 256 1 -caustic 1000000 > /dev/null 2>&1
  echo "static occ $r: $(python profiles/launch_summary.py gpurun_out/pt_$r.csv | grep path_kernel)"
done
