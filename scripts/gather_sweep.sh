mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gather.py -m gpu -q -x > gpurun_out/pytest4.log 2>&1; tail -3 gpurun_out/pytest4.log
GI_CUT_SLACK=40 GI_GATHER_OCC=4 GI_GUESS_FACTOR=1.05 timeout 900 python -m pytest tests/test_gpu_gather.py -m gpu -q -x -k "synthetic or ties or fixture" > gpurun_out/pytest4b.log 2>&1; tail -2 gpurun_out/pytest4b.log
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --aa 1 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 gathers/s %.1fM gather_ms %.1f' % (d['gathers_per_s']/1e6, d['gather_ms']))"; }
export GI_TILE_LEVELS=8
for occ in 3 4; do for g in 1.1 1.25 1.5; do for sl in 64 128 512; do
  GI_GATHER_OCC=$occ GI_GUESS_FACTOR=$g GI_CUT_SLACK=$sl run "occ $occ guess $g slack $sl"
done; done; done
