#!/bin/bash
O=gpurun_out; mkdir -p $O; T=${1:-r02e}
timeout 900 python -m pytest tests/test_gpu_gather.py tests/test_gpu_photons.py tests/test_gpu_cache.py tests/test_zz_no_alloc.py tests/test_gpu_multi.py -x -q 2>&1 | tail -2
timeout 900 python bench.py --no-cpu-baseline --frames small > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err; echo "bench rc=$?"; tail -3 $O/${T}_bench_n1.err
python - <<'PY'
import json,sys
d=json.loads(open('gpurun_out/'+sys.argv[1]+'_bench_n1.json').read().strip().splitlines()[-1]) if len(sys.argv)>1 else None
PY
