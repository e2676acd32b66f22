"""CPU suite: the oracle restatement of the irradiance cache (oracle/restate_cache.cpp) against the reference's own
pre-pass (photonmap.cpp:380-411 + EstimateIrradiance) and EstimateCachedRadiance, run on reference-traced cornell maps
by `ref_harness cache` (tests/golden/cache.npz, made by tests/golden/make_cache_golden.py).
Bar: pre-pass RGBE bytes equal up to one mantissa step where the neighbour sum rounds differently (summation order);
identical chosen photon; radiance within 1e-12 relative."""
import os

import numpy as np
import pytest

from gi_testlib import GOLDEN, GI_QUERY, read_photons, synthetic_photons

G = np.load(os.path.join(GOLDEN, "cache.npz"))
CASES = sorted(k[:-6] for k in G.files if k.endswith("__meta"))
QUERIES = np.load(os.path.join(GOLDEN, "cornell.gather.npz"))["queries"]


def rgbe_close(a, b):
    """same exponent and mantissas within 1 step, for at least 99.9 % identical entries"""
    a = a.astype(int); b = b.astype(int)
    same = (a == b).all(1)
    near = (a[:, 3] == b[:, 3]) & (np.abs(a[:, :3] - b[:, :3]).max(1) <= 1)
    return same.mean() >= 0.999 and (same | near).all()


def side_ok(ph, q):
    """photon_utils.cpp:186-189 with the direction table of :253-272"""
    d = ph["direction"].astype(int)
    phi, theta = d // 256, d % 256
    tp = phi * (2 * np.pi) / 255.0 - np.pi; tt = theta * np.pi / 255.0
    v = np.stack([np.sin(tt) * np.cos(tp), np.sin(tt) * np.sin(tp), np.cos(tt)], 1)
    perp = v @ q["n"]
    return ~(((q["cos_theta"] < 0) & (perp < 0)) | ((q["cos_theta"] > 0) & (perp > 0)))


@pytest.mark.parametrize("case", CASES)
def test_oracle_cache_matches_reference(oracle, case):
    mp, k, r = G[case + "__meta"]; k = int(k); r = float(r)
    ph = read_photons(os.path.join(GOLDEN, f"cornell.{mp}.gipm"))
    rgbe = oracle.irradiance_prepass(ph, k, r, threads=4)
    ref = G[case + "__rgbe"]
    assert rgbe_close(rgbe, ref), (case, (rgbe != ref).any(1).sum())
    assert (ref != ph["rgbe"]).any(1).mean() > 0.9          # the pre-pass did change the colours
    conv = ph.copy(); conv["rgbe"] = ref
    idx, rgb = oracle.cached_radiance(conv, QUERIES, r, threads=4)
    assert np.array_equal(idx, G[case + "__idx"])
    assert np.allclose(rgb, G[case + "__rgb"], rtol=1e-12, atol=1e-300)


def test_oracle_cache_edge_cases(oracle):
    q = np.zeros(4, dtype=GI_QUERY)
    q["n"] = (0, 0, 1); q["bounce"] = (0, 0, 1); q["cos_theta"] = 1.0; q["kd"] = 1.0
    q["p"] = [(0, 0, 1), (5, 5, 5), (0.5, 0.5, 1), (0.25, -0.5, 1)]
    ph = synthetic_photons(400, 3)
    # empty map / nothing in range: black, index -1 (the reference would read through NULL)
    idx, rgb = oracle.cached_radiance(ph[:0], q, 1.0)
    assert (idx == -1).all() and (rgb == 0).all()
    idx, rgb = oracle.cached_radiance(ph, q, 0.3)
    assert idx[1] == -1 and (rgb[1] == 0).all()
    # a photon closer than 1e-6 is never returned (first FindClosest has min_distance = 1e-6)
    near = ph.copy(); near["position"][7] = (0.5, 0.5, 1.0 - 5e-7)
    idx, _ = oracle.cached_radiance(near, q, 1.0)
    assert idx[2] != 7
    # wrong-sided photons are skipped: the result is the nearest photon whose direction passes the side test
    idx, _ = oracle.cached_radiance(ph, q, 3.0)
    for i in (0, 2, 3):
        d2 = ((ph["position"] - q["p"][i]) ** 2).sum(1)
        ok = side_ok(ph, q[i]) & (d2 >= 1e-12) & (d2 <= 9.0)
        assert idx[i] == np.flatnonzero(ok)[np.argmin(d2[ok])]
    # pre-pass: k = 1 finds the photon itself (d = 0): colour' = colour + colour / (pi * 1e-6)
    rgbe = oracle.irradiance_prepass(ph[:1], 1, 1.0)
    assert rgbe[0, 3] > ph["rgbe"][0, 3]
