"""The compact host-buffer entry points (gi_intersect_batch_compact: 16-byte hits; gi_gather_batch_compact: 80-byte
queries with a material index) against the reference-shaped ones on the same inputs: identical hit ids and t, identical
photon counts, radiance to 1e-12 relative (the bounce direction is derived on the device instead of on the host)."""
import os
import numpy as np
import pytest

import gi_b200
from gi_testlib import GOLDEN, GI_QUERY, random_rays, synthetic_photons

pytestmark = pytest.mark.gpu


def test_compact_variants_equal_the_reference_shaped_calls():
    scene = gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc"))
    v = scene.view
    ctx = gi_b200.Context(0)
    try:
        ctx.upload_scene(scene)
        rays = random_rays(60000, list(v.bbox_min), list(v.bbox_max), 5, scale=1.2)
        full = ctx.intersect(rays)
        comp = ctx.intersect_compact(rays)
        hit = full["hit"] == 1
        assert hit.sum() > 20000 and (~hit).sum() > 1000
        assert np.array_equal(comp["prim"][hit], full["prim"][hit]) and np.array_equal(comp["shape"][hit], full["shape"][hit])
        assert np.array_equal(comp["t"][hit], full["t"][hit])                     # bit-equal
        assert (comp["prim"][~hit] == -1).all()
        # queries at the hits, as RayTrace builds them (raytracer.cpp:138-149): the full records on the host ...
        h = full[hit]; d = rays[hit][:, 3:6]
        mats, els = scene.materials(), scene.elements()
        mat = els["material"][h["element"]].astype(np.int32)
        kd, ks, sh = mats["kd"], mats["ks"], mats["shininess"]
        n = h["n"].copy(); cos = -(n * d).sum(1)
        nn = np.where((cos < 0)[:, None], -n, n); cc = np.abs(cos)                # ReflectiveBounce, graphics_utils.cpp:104-117
        b = d + nn * cc[:, None] * 2.0
        b /= np.sqrt((b[:, 0] * b[:, 0] + b[:, 1] * b[:, 1] + b[:, 2] * b[:, 2]))[:, None]
        q = np.zeros(len(h), dtype=GI_QUERY)
        q["p"] = h["p"]; q["n"] = n; q["bounce"] = b; q["cos_theta"] = cos
        q["kd"] = kd[mat]; q["ks"] = ks[mat]; q["shininess"] = sh[mat]
        qc = np.zeros(len(h), dtype=gi_b200.QUERY_COMPACT)
        qc["p"] = h["p"]; qc["n"] = n; qc["view"] = d; qc["material"] = mat
        # ... on a map lying on the same surfaces
        ph = synthetic_photons(len(h), 3); ph["position"] = h["p"] + np.random.default_rng(2).normal(scale=0.02, size=(len(h), 3))
        ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
        ctx.build_maps()
        for k, r, filt in ((50, 0.4, 0), (25, 0.2, 1), (10, 0.3, 2)):
            _, cnt, rgb = ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt, want_idx=False)
            cnt_c, rgb_c = ctx.gather_compact(gi_b200.GI_MAP_CAUSTIC, qc, k, r, filt)
            assert np.array_equal(cnt, cnt_c) and (cnt == k).sum() > len(h) // 2
            scale = np.abs(rgb).max(axis=1, keepdims=True)
            assert (np.abs(rgb - rgb_c) <= 1e-12 * scale + 1e-300).all()
    finally:
        ctx.close()
