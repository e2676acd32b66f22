#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ from the REFERENCE ITSELF (run in the build container only).

Needs /root/reference (scene inputs) and oracle/_ref/ref_harness (the unmodified reference sources compiled in
place by oracle/Makefile).  Nothing here is read at test time except the files it writes:

  <scene>.gisc        flattened scene written by OUR loader (global-illumination_b200/bin/gi_scene_tool) -- checked
                      against the reference's own scene graph by tests/test_loader.py via the text dumps
  <scene>.dump.txt.gz `ref_harness dump_scene` text dump of the reference's R3Scene graph
  <scene>.isect.npz   rays (camera grid + uniform random + secondary rays leaving hit points) and the hit records
                      returned by the reference's R3Scene::Intersects for them
  cornell.*.gipm      photon maps traced by the reference's MapPhotons (unseeded RNG: these files ARE the fixture)
  cornell.gather.npz  queries + kNN index sets + radiance from the reference's R3Kdtree / EstimateRadiance
"""
import gzip
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa

SCENES = ["cornell", "jensen", "teapot", "stilllife", "transform", "cylinder", "stack", "distributed", "box", "sphere",
          "lines", "ico", "art3", "glass_cube", "specular", "softshadow"]
TMP = "/tmp/gi_golden"


def make_rays(scn, gisc, name):
    ref_harness("camrays", scn, 32, 32, f"{TMP}/{name}.cam.bin")
    cam = read_rays(f"{TMP}/{name}.cam.bin")
    tail = np.frombuffer(open(gisc, "rb").read()[-128:], dtype="<f8")
    rnd = random_rays(1500, tail[6:9], tail[9:12], seed=11)
    first = np.concatenate([cam, rnd])
    write_rays(f"{TMP}/{name}.r1.bin", first)
    ref_harness("isect", scn, f"{TMP}/{name}.r1.bin", f"{TMP}/{name}.h1.bin", 8)
    h1 = read_ref_hits(f"{TMP}/{name}.h1.bin")
    m = h1["hit"] == 1
    rng = np.random.default_rng(5)
    d2 = rng.normal(size=(m.sum(), 3)); d2 /= np.sqrt((d2 * d2).sum(1))[:, None]
    d2 = np.where(((d2 * h1["n"][m]).sum(1) < 0)[:, None], -d2, d2)
    # origins ON the surface, 1e-6 and 0.5e-6 off it: exercises the reference's negative-t rules
    o2 = h1["p"][m] + 1e-6 * d2 * rng.choice([0.0, 1.0, 0.5], size=(m.sum(), 1))
    sec = np.concatenate([o2, d2], axis=1)[:1200]
    return np.concatenate([first, sec])


def main():
    os.makedirs(TMP, exist_ok=True)
    for name in SCENES:
        scn = f"{REF_INPUT}/{name}.scn"
        gisc = f"{HERE}/{name}.gisc"
        flatten_scene(scn, gisc)
        ref_harness("dump_scene", scn, f"{TMP}/{name}.dump.txt")
        with open(f"{TMP}/{name}.dump.txt", "rb") as f, gzip.GzipFile(f"{HERE}/{name}.dump.txt.gz", "wb", mtime=0) as g:
            g.write(f.read())
        rays = make_rays(scn, gisc, name)
        write_rays(f"{TMP}/{name}.rays.bin", rays)
        ref_harness("isect", scn, f"{TMP}/{name}.rays.bin", f"{TMP}/{name}.hits.bin", 8)
        hits = read_ref_hits(f"{TMP}/{name}.hits.bin")
        np.savez_compressed(f"{HERE}/{name}.isect.npz", rays=rays, hits=hits)
        print(name, "rays", len(rays), "hits", int(hits["hit"].sum()))
    # photon maps + gather
    scn = f"{REF_INPUT}/cornell.scn"
    print(ref_harness("photons", scn, f"{HERE}/cornell", "-global", 2176, "-caustic", 20000, "-threads", 8).stdout)
    d = np.load(f"{HERE}/cornell.isect.npz")
    sel = np.nonzero(d["hits"]["hit"][:1024] == 1)[0][::3][:300]
    hits = np.zeros(len(sel), dtype=GI_HIT)
    for f in ("hit", "t", "p", "n"):
        hits[f] = d["hits"][f][sel]
    q = queries_from_hits(d["rays"][sel], hits)
    write_queries(f"{TMP}/q.bin", q)
    out = {"queries": q}
    for mp, k, r in (("caustic", 225, 0.225), ("global", 50, 2.5), ("caustic", 40, 0.05)):
        for filt in (0, 1, 2):
            ref_harness("gather", f"{HERE}/cornell.{mp}.gipm", f"{TMP}/q.bin", f"{TMP}/g.bin", k, r, filt, 8)
            cnt, rgb, idx = read_ref_gather(f"{TMP}/g.bin")
            key = f"{mp}_{k}_{filt}"
            out[key + "_count"] = cnt; out[key + "_rgb"] = rgb
            if filt == 0:
                out[f"{mp}_{k}_idx"] = idx
            out[key + "_r"] = np.array([r])
    np.savez_compressed(f"{HERE}/cornell.gather.npz", **out)
    print("gather fixture queries", len(q))


if __name__ == "__main__":
    main()
