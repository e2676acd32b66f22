#!/bin/bash
# Round-2 profile captures (one GPU).  Every ncu command runs only after the same command exited 0 without ncu.
set -x
O=gpurun_out; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.log 2>&1; tail -2 $O/r02_smoke.log
# launch list of the bench command (small step so that ncu's serialised replay stays short)
python bench.py --steps 2 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --frames none --no-extras > $O/b_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r02_launches_bench_small.csv \
    python bench.py --steps 2 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --frames none --no-extras > $O/ncu_b.log 2>&1
# launch list of a whole-program frame
REPS=1 python scripts/render_perf.py cornell 256 1 > $O/f_small.log 2>&1 && \
REPS=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/r02_launches_frame_cornell256aa1.csv \
    python scripts/render_perf.py cornell 256 1 > $O/ncu_f.log 2>&1
# ncu --set full of the dominant kernel of the bench step (gather_kernel, k = 225, 1e7 photons, 2^20 queries)
ncu --set full --clock-control none --import-source on -k regex:gather_kernel -s 1 -c 1 -f -o $O/r02_gather_kernel \
    python bench.py --steps 1 --warmup 1 --width 512 --height 512 --aa 1 --no-cpu-baseline --frames none --no-extras > $O/ncu_g.log 2>&1
# the small-map kernels at the bench section's size (2^23 queries)
python scripts/gather_small.py 2900 > $O/r02_gather_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:small -s 4 -c 4 --csv --log-file $O/r02_small_map_kernels.csv python scripts/gather_small.py 2900 > $O/ncu_s.log 2>&1
cat $O/r02_gather_small.log
# ncu --set full of the first closest-hit pass of a path batch
REPS=1 ncu --set full --clock-control none --import-source on -k regex:wave_hit_kernel -s 0 -c 1 -f -o $O/r02_wave_hit \
   python scripts/render_perf.py cornell 256 1 > $O/ncu_h.log 2>&1
