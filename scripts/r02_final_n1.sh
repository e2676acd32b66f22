#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python bench.py --frames all > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; echo "bench rc=$?"; tail -3 $O/r02_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench_reference_arm.json 2> $O/r02_bench_ref.err; echo "ref rc=$?"
