#!/bin/bash
# render-time sensitivity to the gather tunables (GPU box)
B=global-illumination_b200/bin/photonmap
run() { echo -n "$1: "; env $1 $B tests/golden/${2:-cornell}.gisc gpurun_out/c.png -resolution 256 256 -aa 1 -caustic 1000000 -v | grep -A1 "Rendered image" | grep Time; }
for s in cornell jensen; do
run GI_SORT_REQUESTS=0 $s
run GI_SORT_REQUESTS=1 $s
done
