#!/bin/bash
# wavefront path stage vs the round-1 thread-per-path kernel, same box
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_cli.py -x -q > $O/r02b_tests.log 2>&1; tail -3 $O/r02b_tests.log
for rep in 1 2; do
GI_PATH_MEGAKERNEL=1 REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed 's/^/mega  /'
REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed 's/^/wave4 /'
done
GI_WAVE_PASSES=2 REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed 's/^/wave2 /'
GI_WAVE_PASSES=8 REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed 's/^/wave8 /'
GI_PATH_MEGAKERNEL=1 REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1 | sed 's/^/mega  /'
REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1 | sed 's/^/wave4 /'
