#!/usr/bin/env python
"""Generate tests/golden/cache.npz from the REFERENCE ITSELF (run in the build container only).

`oracle/_ref/ref_harness cache` runs the irradiance-cache pre-pass of MapPhotons (src/photonmap.cpp:380-411, around
the reference's own EstimateIrradiance) on the reference-traced cornell global map and then the reference's own
EstimateCachedRadiance (src/utils/photon_utils.cpp:164-205) at the queries of cornell.gather.npz.

  <case>__rgbe    colours of the photons after the pre-pass (n x 4 bytes)
  <case>__rgb     EstimateCachedRadiance per query (n x 3)
  <case>__idx     the photon it settled on (-1: none within r; the reference would crash there, the harness skips)
  cases: (k, r) = GLOBAL_ESTIMATE_SIZE / _DIST; the caustic map (20 k photons) is used as a second, denser map.
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa

TMP = "/tmp/gi_golden_cache"
CASES = {"global_k50_r2.5": ("global", 50, 2.5), "global_k30_r1.5": ("global", 30, 1.5), "global_k8_r0.2": ("global", 8, 0.2),
         "caustic_k50_r0.5": ("caustic", 50, 0.5), "caustic_k100_r0.05": ("caustic", 100, 0.05)}


def main():
    os.makedirs(TMP, exist_ok=True)
    q = np.load(f"{HERE}/cornell.gather.npz")["queries"]
    write_queries(f"{TMP}/q.bin", q)
    out = {}
    for name, (mp, k, r) in CASES.items():
        ref_harness("cache", f"{HERE}/cornell.{mp}.gipm", f"{TMP}/q.bin", f"{TMP}/c.bin", k, r, f"{TMP}/c.gipm")
        cnt, rgb, idx = read_ref_gather(f"{TMP}/c.bin")
        ph = read_photons(f"{TMP}/c.gipm")
        out[name + "__rgbe"] = ph["rgbe"].copy()
        out[name + "__rgb"] = rgb
        out[name + "__idx"] = idx[:, 0]
        out[name + "__meta"] = np.array([mp, str(k), str(r)])
        print(name, "photons", len(ph), "queries", len(q), "no photon in range:", int((cnt < 0).sum()), "mean rgb", rgb.mean(0))
    np.savez_compressed(f"{HERE}/cache.npz", **out)


if __name__ == "__main__":
    main()
