#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_isect.py tests/test_gpu_render.py tests/test_gpu_photons.py -x -q 2>&1 | tail -2
REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1
REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1
timeout 300 python bench.py --no-cpu-baseline --frames none --steps 3 --warmup 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('isect_ms %.3f  value %.1f  mesh cam %.0f rnd %.0f Mrays/s  small-map %.0f M/s' % (d['intersect_ms'], d['value'], d['intersect_mesh']['camera']['mrays_per_s'], d['intersect_mesh']['random']['mrays_per_s'], d['gathers_small_map']['gathers_per_s']/1e6))
"
