#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_gather.py tests/test_gpu_cache.py tests/test_gpu_render.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
for v in 0 1; do GI_GATHER_SMALL=$v timeout 120 python scripts/gather_small.py 2>&1 | tail -1 | sed "s/^/small=$v /"; done
for v in 0 1; do GI_GATHER_SMALL=$v REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed "s/^/small=$v /"; done
for v in 0 1; do GI_GATHER_SMALL=$v REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1 | sed "s/^/small=$v /"; done
