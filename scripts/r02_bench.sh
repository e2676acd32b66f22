#!/bin/bash
# bench (N = 1), reference arm, small-map gather under ncu
O=gpurun_out; mkdir -p $O; T=${1:-r02d}
timeout 900 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err; echo "bench rc=$?"; tail -3 $O/${T}_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference_arm.json 2> $O/${T}_bench_ref.err; echo "ref rc=$?"
timeout 120 python scripts/gather_small.py > $O/${T}_gather_small.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gather_kernel -s 2 -c 1 -f -o $O/${T}_gather_small \
   python scripts/gather_small.py > $O/${T}_gather_small_ncu.log 2>&1
cat $O/${T}_gather_small.log
