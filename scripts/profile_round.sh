#!/bin/bash
# Round profile capture (GPU box, 1 GPU).  Every ncu command runs only after the same command exited 0 without ncu.
# Outputs under gpurun_out/ ; the summaries that are judged get copied to profiles/ by hand.
set -x
R=${1:-r01}
O=gpurun_out
mkdir -p $O
python bench.py --steps 5 --warmup 3 > $O/${R}_bench_n1.json 2> $O/${R}_bench_n1.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/${R}_bench_reference_arm.json 2> $O/${R}_bench_ref.err
# launch list of the bench command (small step so that ncu's serialised replay stays short)
python bench.py --steps 2 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --no-frame > $O/b_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${R}_launches_bench_small.csv \
    python bench.py --steps 2 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --no-frame > $O/ncu_b.log 2>&1
# launch list of a whole-program frame
REPS=1 python scripts/render_perf.py cornell 256 1 > $O/f_small.log 2>&1 && \
REPS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/${R}_launches_frame_cornell256aa1.csv \
    python scripts/render_perf.py cornell 256 1 > $O/ncu_f.log 2>&1
