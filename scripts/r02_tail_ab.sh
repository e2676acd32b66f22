#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_cli.py tests/test_gpu_cache.py -x -q 2>&1 | tail -2
for v in 1 0; do
 if [ $v = 1 ]; then export GI_NO_TAIL_OVERLAP=1; else unset GI_NO_TAIL_OVERLAP; fi
 REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed "s/^/no_overlap=$v /"
 REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1 | sed "s/^/no_overlap=$v /"
done
unset GI_NO_TAIL_OVERLAP
REPS=1 timeout 300 python scripts/render_perf.py cornell 1024 2 2>&1 | tail -1
