#!/usr/bin/env python
"""jensen.scn photon maps + gathers from the REFERENCE ITSELF (build container only).

  jensen.gather.npz   global_photons / caustic_photons: maps traced by the reference's MapPhotons
                      (`ref_harness photons`, -global 2176 -caustic 150000; unseeded RNG: the arrays ARE the fixture),
                      queries at the reference's primary hits of a 48 x 48 camera grid + on photons of the caustic,
                      and kNN index sets / counts / radiance from the reference's R3Kdtree::FindClosestQuick +
                      EstimateRadiance (`ref_harness gather`) for k = 225 / 50, three filters.
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa
from make_golden import TMP


def main():
    os.makedirs(TMP, exist_ok=True)
    scn = f"{REF_INPUT}/jensen.scn"
    print(ref_harness("photons", scn, f"{TMP}/jensen", "-global", 2176, "-caustic", 150000, "-threads", 8).stdout)
    gph = read_photons(f"{TMP}/jensen.global.gipm"); cph = read_photons(f"{TMP}/jensen.caustic.gipm")
    ref_harness("camrays", scn, 48, 48, f"{TMP}/jensen.cam.bin")
    rays = read_rays(f"{TMP}/jensen.cam.bin")
    ref_harness("isect", scn, f"{TMP}/jensen.cam.bin", f"{TMP}/jensen.cam.hits", 8)
    h = read_ref_hits(f"{TMP}/jensen.cam.hits")
    hits = np.zeros(len(h), dtype=GI_HIT)
    for f in ("hit", "t", "p", "n"):
        hits[f] = h[f]
    q = queries_from_hits(rays, hits)[::2][:900]
    rng = np.random.default_rng(3)
    extra = q[rng.integers(0, len(q), 300)].copy()
    extra["p"] = cph["position"][rng.integers(0, len(cph), 300)] + rng.normal(scale=1e-4, size=(300, 3))
    q = np.concatenate([q, extra])
    write_queries(f"{TMP}/jq.bin", q)
    out = {"queries": q.view(np.uint8), "global_photons": gph.view(np.uint8), "caustic_photons": cph.view(np.uint8)}
    for mp, k, r, filters in (("caustic", 225, 0.225, (0, 1, 2)), ("caustic", 50, 0.1, (0,)), ("global", 50, 2.5, (0, 2))):
        for filt in filters:
            ref_harness("gather", f"{TMP}/jensen.{mp}.gipm", f"{TMP}/jq.bin", f"{TMP}/jg.bin", k, r, filt, 8)
            cnt, rgb, idx = read_ref_gather(f"{TMP}/jg.bin")
            key = f"{mp}_{k}_{filt}"
            out[key + "_count"] = cnt; out[key + "_rgb"] = rgb; out[key + "_r"] = np.array([r])
            out[f"{mp}_{k}_idx"] = idx
            print(key, "full", int((cnt == k).sum()), "of", len(q))
    np.savez_compressed(f"{HERE}/jensen.gather.npz", **out)


if __name__ == "__main__":
    main()
