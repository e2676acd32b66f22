#!/usr/bin/env python
"""Round-2 fixtures, made by the REFERENCE ITSELF (run in the build container only: needs /root/reference and
oracle/_ref/ref_harness).  Same products per scene as make_golden.py (flattened scene by OUR loader, the reference's
scene-graph dump, rays + the reference's R3Scene::Intersects hit records):

  circle           tests/golden/scn/circle.scn, hand-written here: R3Circle primitives (R3Isect.cpp:837-879), which no
                   input scene of the reference uses
  distributed_lit  tests/golden/scn/distributed_lit.scn: the lit bench variant of input/distributed.scn (configs[4])
  art2             input/art2.scn: the glass teapot (teapot.off under a scaled group, material 6: kt .92, ir 2.4) in a
                   box with a rect light -- the "refraction + Fresnel" teapot of configs[2]
  violin           input/violin.scn (configs[3]: 1 box + a 104 300-triangle mesh); the flattened scene is stored
                   gzip'd (9.2 MB -> 1.5 MB), gi_testlib.golden_gisc() unpacks it
"""
import gzip
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa
from make_golden import make_rays, TMP

CASES = [("circle", f"{HERE}/scn/circle.scn", False), ("distributed_lit", f"{HERE}/scn/distributed_lit.scn", False),
         ("art2", f"{REF_INPUT}/art2.scn", False), ("violin", f"{REF_INPUT}/violin.scn", True)]


def main():
    os.makedirs(TMP, exist_ok=True)
    for name, scn, packed in CASES:
        gisc = f"{TMP}/{name}.gisc" if packed else f"{HERE}/{name}.gisc"
        flatten_scene(scn, gisc)
        if packed:
            with open(gisc, "rb") as f, gzip.GzipFile(f"{HERE}/{name}.gisc.gz", "wb", mtime=0) as g:
                g.write(f.read())
        else:
            ref_harness("dump_scene", scn, f"{TMP}/{name}.dump.txt")
            with open(f"{TMP}/{name}.dump.txt", "rb") as f, gzip.GzipFile(f"{HERE}/{name}.dump.txt.gz", "wb", mtime=0) as g:
                g.write(f.read())
        rays = make_rays(scn, gisc, name)
        write_rays(f"{TMP}/{name}.rays.bin", rays)
        ref_harness("isect", scn, f"{TMP}/{name}.rays.bin", f"{TMP}/{name}.hits.bin", os.cpu_count() or 8)
        hits = read_ref_hits(f"{TMP}/{name}.hits.bin")
        np.savez_compressed(f"{HERE}/{name}.isect.npz", rays=rays, hits=hits)
        print(name, "rays", len(rays), "hits", int(hits["hit"].sum()))


if __name__ == "__main__":
    main()
