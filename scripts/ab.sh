#!/bin/bash
# A/B of two prebuilt libraries (build_ab/libA.so, libB.so) on the same box: bench gather rate, alternating
for rep in 1 2; do for v in ${VARIANTS:-A B}; do
cp build_ab/lib$v.so global-illumination_b200/libgi_b200.so
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-frame 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$v gathers/s %.4g  gather_ms %.2f value %.1f e2e %.1f clocks %s' % (d['gathers_per_s'], d['gather_ms'], d['value'], d['e2e']['value'], d['clocks']))
"
done; done
# frame timing and gather parity with the last variant left in place
[ -n "$FRAME" ] && for v in ${VARIANTS:-A B}; do cp build_ab/lib$v.so global-illumination_b200/libgi_b200.so; echo "== $v"; REPS=2 timeout 120 python scripts/render_perf.py cornell 512 1; done
[ -n "$TESTS" ] && (timeout 300 python -m pytest tests/test_gpu_gather.py tests/test_gpu_cache.py -x -q 2>&1 | tail -3)
true
