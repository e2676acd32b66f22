#!/bin/bash
# A/B of prebuilt libraries (build_ab/lib<V>.so) on the photon pass
for v in $VARIANTS; do cp build_ab/lib$v.so global-illumination_b200/libgi_b200.so; echo "== $v"; timeout 200 python scripts/photon_perf.py 2>&1 | tail -2; done
