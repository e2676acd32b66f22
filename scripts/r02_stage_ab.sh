#!/bin/bash
GI_GATHER_STAGE=1 timeout 900 python -m pytest tests/test_gpu_gather.py tests/test_gpu_cache.py -x -q 2>&1 | tail -2
for rep in 1 2; do for v in 0 1; do
GI_GATHER_STAGE=$v timeout 300 python bench.py --no-cpu-baseline --frames none --no-extras --steps 3 --warmup 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('stage=$v gather_ms %.2f gathers/s %.4g value %.1f' % (d['gather_ms'], d['gathers_per_s'], d['value']))
"
done; done
for v in 0 1; do GI_GATHER_STAGE=$v REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1 | sed "s/^/stage=$v /"; done
for v in 0 1; do GI_GATHER_STAGE=$v timeout 120 python scripts/gather_small.py 2>&1 | tail -1 | sed "s/^/stage=$v /"; done
