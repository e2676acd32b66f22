#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_gather.py tests/test_gpu_cache.py tests/test_gpu_photons.py -x -q 2>&1 | tail -3
for v in 1 0; do
GI_KD_SORT_BUILD=$v timeout 300 python bench.py --no-cpu-baseline --frames none --no-extras --steps 3 --warmup 2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('sort_build=$v value %.1f ms/step %.1f stages %s' % (d['value'], d['ms_per_step'], {k:round(x,2) for k,x in d['stage_ms'].items()}))
"
done
