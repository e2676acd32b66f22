#!/bin/bash
# full GPU suite + frame timings
O=gpurun_out; mkdir -p $O; T=${1:-r02c}
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${T}_gputests.log 2>&1; tail -3 $O/${T}_gputests.log
REPS=2 timeout 300 python scripts/render_perf.py cornell 512 1 2>&1 | tail -1
REPS=2 timeout 300 python scripts/render_perf.py jensen 512 1 2>&1 | tail -1
REPS=1 timeout 300 python scripts/render_perf.py cornell 1024 2 2>&1 | tail -1
