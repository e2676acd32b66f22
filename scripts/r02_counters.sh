#!/bin/bash
# instrumented gather (GI_GATHER_DEBUG=1): per-query walk counters of the bench step's launch and of the small-map kernels
O=gpurun_out; mkdir -p $O
GI_GATHER_DEBUG=1 timeout 600 python bench.py --no-cpu-baseline --frames none --steps 1 --warmup 0 2>&1 >/dev/null | grep "^\[gather" | sort | uniq -c | sort -rn | head -6 > $O/r02_gather_counters.txt
cat $O/r02_gather_counters.txt
