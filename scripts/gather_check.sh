#!/bin/bash
# gather parity tests + bench gather rate (+ optional ncu capture: TAG given); extra env settings are swept after
(python -m pytest tests/test_gpu_gather.py tests/test_gpu_cache.py -x -q 2>&1 | tail -8)
run() {
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-frame 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1 gathers/s %.4g  gather_ms %.2f  kd_build_ms %.1f  value %.1f  e2e %.1f' % (d['gathers_per_s'], d['gather_ms'], d['kd_build_ms'], d['value'], d['e2e']['value']))
    else: print(l.rstrip())
"
}
run default
for s in $SWEEP; do env $s bash -c "$(declare -f run); run $s"; done
[ -n "$1" ] && scripts/ncu_gather.sh $1
