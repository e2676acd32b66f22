"""The irradiance cache (-cache) on the GPU: gi_build_maps' pre-pass and gi_cached_radiance_batch against the
reference (tests/golden/cache.npz, made by `ref_harness cache`) and against the oracle restatement on synthetic maps."""
import os

import numpy as np
import pytest

import gi_b200
from gi_testlib import GOLDEN, GI_QUERY, read_photons, synthetic_photons
from test_oracle_cache import rgbe_close

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(GOLDEN, "cache.npz"))
CASES = sorted(k[:-6] for k in G.files if k.endswith("__meta"))
QUERIES = np.load(os.path.join(GOLDEN, "cornell.gather.npz"))["queries"]


def cache_params(k, r):
    p = gi_b200.default_params()
    p.irradiance_cache = 1; p.global_estimate_size = k; p.global_estimate_dist = r
    return p


def build(ctx, ph, k, r):
    ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, np.zeros(0, dtype=gi_b200.PHOTON))
    ctx.load_photons(gi_b200.GI_MAP_GLOBAL, ph)
    ctx.build_maps(cache_params(k, r))
    return ctx.get_photons(gi_b200.GI_MAP_GLOBAL)


@pytest.mark.parametrize("case", CASES)
def test_cache_matches_reference(gpu_ctx, case):
    mp, k, r = G[case + "__meta"]; k = int(k); r = float(r)
    ph = read_photons(os.path.join(GOLDEN, f"cornell.{mp}.gipm"))
    gpu_ctx.upload_scene(gi_b200.HostScene(os.path.join(GOLDEN, "cornell.gisc")))
    conv = build(gpu_ctx, ph, k, r)
    assert np.array_equal(conv["position"], ph["position"]) and np.array_equal(conv["direction"], ph["direction"])
    assert rgbe_close(conv["rgbe"], G[case + "__rgbe"]), (conv["rgbe"] != G[case + "__rgbe"]).any(1).sum()
    with pytest.raises(gi_b200.GiError, match="cached irradiance"):   # such a map cannot be rendered without -cache
        gpu_ctx.render_float(gi_b200.parse_flags(["-photon_viz"])[0], 8, 8, 0)
    rgb = gpu_ctx.cached_radiance(QUERIES, r)
    ref, idx = G[case + "__rgb"], G[case + "__idx"]
    assert (rgb[idx < 0] == 0).all()                  # nothing in range: black (the reference would crash)
    # where the chosen photon's converted colour is byte-identical to the reference's, so is the radiance
    same = np.zeros(len(idx), dtype=bool)
    same[idx >= 0] = (conv["rgbe"][idx[idx >= 0]] == G[case + "__rgbe"][idx[idx >= 0]]).all(1)
    assert same[idx >= 0].mean() > 0.99
    assert np.allclose(rgb[same], ref[same], rtol=1e-12, atol=1e-300)
    assert (ref[same].sum(1) > 0).sum() > 100


def test_cache_needs_the_prepass(gpu_ctx):
    ph = read_photons(os.path.join(GOLDEN, "cornell.global.gipm"))
    gpu_ctx.load_photons(gi_b200.GI_MAP_GLOBAL, ph)
    gpu_ctx.build_maps(gi_b200.default_params())
    with pytest.raises(gi_b200.GiError, match="irradiance cache"):
        gpu_ctx.cached_radiance(QUERIES, 2.5)
    # building again with the flag set runs the pre-pass on the already built tree, once
    p = cache_params(50, 2.5)
    gpu_ctx.build_maps(p)
    a = gpu_ctx.get_photons(gi_b200.GI_MAP_GLOBAL)
    gpu_ctx.build_maps(p)
    b = gpu_ctx.get_photons(gi_b200.GI_MAP_GLOBAL)
    assert np.array_equal(a, b) and (a["rgbe"] != ph["rgbe"]).any(1).mean() > 0.9


@pytest.mark.parametrize("n,k,r,seed", [(3000, 20, 0.3, 1), (50000, 50, 0.1, 2), (200, 64, 5.0, 3), (1, 4, 1.0, 4)])
def test_cache_matches_oracle_on_synthetic_maps(gpu_ctx, oracle, n, k, r, seed):
    ph = synthetic_photons(n, seed)
    if n > 1000:
        ph["position"][n // 2:n // 2 + 50] = ph["position"][:50]      # exact duplicates: ties by lower index
    conv = build(gpu_ctx, ph, k, r)
    want = oracle.irradiance_prepass(ph, k, r)
    assert rgbe_close(conv["rgbe"], want)
    rng = np.random.default_rng(seed)
    m = 4000
    q = np.zeros(m, dtype=GI_QUERY)
    src = ph["position"][rng.integers(0, n, m)]
    q["p"] = src + rng.normal(0, 0.02, (m, 3)) * (rng.random((m, 1)) < 0.8)    # a fifth of the queries sit ON a photon
    nrm = rng.normal(size=(m, 3)); q["n"] = nrm / np.linalg.norm(nrm, axis=1, keepdims=True)
    b = rng.normal(size=(m, 3)); q["bounce"] = b / np.linalg.norm(b, axis=1, keepdims=True)
    q["cos_theta"] = rng.uniform(-1, 1, m); q["kd"] = rng.random((m, 3)); q["ks"] = rng.random((m, 3)) * (rng.random((m, 1)) < 0.5)
    q["shininess"] = rng.uniform(1, 50, m)
    for radius in (r, 4 * r):
        rgb = gpu_ctx.cached_radiance(q, radius)
        idx, ref = oracle.cached_radiance(conv, q, radius)
        assert np.allclose(rgb, ref, rtol=1e-12, atol=1e-300), np.abs(rgb - ref).max()
        assert ((rgb == 0).all(1) >= (idx < 0)).all()
    assert (idx >= 0).mean() > (0.5 if n > 1 else 0.2)      # one photon: the side test alone rejects half


def test_cache_empty_map(gpu_ctx):
    gpu_ctx.load_photons(gi_b200.GI_MAP_GLOBAL, np.zeros(0, dtype=gi_b200.PHOTON))
    gpu_ctx.build_maps(cache_params(50, 2.5))
    rgb = gpu_ctx.cached_radiance(QUERIES[:10], 2.5)
    assert (rgb == 0).all()
