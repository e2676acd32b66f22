#!/bin/bash
# Round-2 baseline on one B200: frame timings, launch list of a frame, ncu --set full of one small-map gather launch and one path_kernel launch
O=gpurun_out; mkdir -p $O
REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 > $O/r02a_frame512.log 2>&1
REPS=1 timeout 300 python scripts/render_perf.py cornell 256 1 > $O/r02a_f_small.log 2>&1 && \
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/r02a_launches_frame_cornell256aa1.csv \
    python scripts/render_perf.py cornell 256 1 > $O/r02a_ncu_f.log 2>&1
REPS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:gather_kernel -s 6 -c 1 -f -o $O/r02a_framegather \
   python scripts/render_perf.py cornell 256 1 > $O/r02a_ncu_g.log 2>&1
REPS=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:path_kernel -s 10 -c 1 -f -o $O/r02a_path \
   python scripts/render_perf.py cornell 256 1 > $O/r02a_ncu_p.log 2>&1
cat $O/r02a_frame512.log
