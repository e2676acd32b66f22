#!/bin/bash
# Times the photonmap program on a few named scenes (GPU box).  Usage: scripts/cli_perf.sh [res] [aa]
RES=${1:-256}; AA=${2:-0}
B=global-illumination_b200/bin/photonmap
mkdir -p gpurun_out
for s in cornell jensen; do
  echo "=== $s ${RES}x${RES} aa $AA defaults"
  time timeout 600 $B tests/golden/$s.gisc gpurun_out/$s.png -resolution $RES $RES -aa $AA -v 2>&1 | grep -v "^\[" | tr '\r' '\n' | grep -v "^$"
done
