"""GPU suite (pytest -m gpu): gi_build_maps + gi_gather_batch through the C ABI against (1) the reference's own
R3Kdtree / EstimateRadiance outputs on reference-traced cornell maps (golden fixture), (2) the brute-force oracle on
synthetic maps, plus size-independent properties at large sizes.  Bar (BASELINE.json): bit-exact kNN index sets (ties
by lower index); radiance within 1e-4 relative (we assert 1e-9: only the summation order differs)."""
import os

import numpy as np
import pytest

import gi_b200
from gi_testlib import ROOT, read_photons, synthetic_photons, GI_QUERY

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(ROOT, "tests", "golden")


def surface_queries(n, seed, kd=(0.5, 0.6, 0.7), ks=(0.2, 0.1, 0.3), shin=10.0):
    rng = np.random.default_rng(seed)
    ph = synthetic_photons(n, seed + 1000)
    q = np.zeros(n, dtype=GI_QUERY)
    q["p"] = ph["position"] + rng.normal(scale=1e-3, size=(n, 3))
    nn = rng.normal(size=(n, 3)); nn /= np.linalg.norm(nn, axis=1)[:, None]
    b = rng.normal(size=(n, 3)); b /= np.linalg.norm(b, axis=1)[:, None]
    q["n"] = nn; q["bounce"] = b; q["cos_theta"] = rng.uniform(-1, 1, n)
    q["kd"] = kd; q["ks"] = ks; q["shininess"] = shin
    return q


def check(ref, got, rtol=1e-9):
    ridx, rcnt, rrgb = ref; idx, cnt, rgb = got
    assert np.array_equal(rcnt, cnt)
    bad = np.nonzero((ridx != idx).any(1))[0]
    assert len(bad) == 0, f"kNN sets differ for queries {bad[:5]}"
    scale = np.abs(rrgb).max(axis=1, keepdims=True)
    assert (np.abs(rgb - rrgb) <= rtol * scale + 1e-300).all()


@pytest.mark.parametrize("mp,k", [("caustic", 225), ("global", 50), ("caustic", 40)])
@pytest.mark.parametrize("filt", [0, 1, 2])
def test_cuda_gather_matches_reference_fixture(gpu_ctx, mp, k, filt):
    fx = np.load(os.path.join(GOLDEN, "cornell.gather.npz"))
    which = gi_b200.GI_MAP_CAUSTIC if mp == "caustic" else gi_b200.GI_MAP_GLOBAL
    gpu_ctx.load_photons_file(which, os.path.join(GOLDEN, f"cornell.{mp}.gipm"))
    gpu_ctx.build_maps()
    key = f"{mp}_{k}_{filt}"
    got = gpu_ctx.gather(which, fx["queries"], k, float(fx[key + "_r"][0]), filt)
    check((fx[f"{mp}_{k}_idx"], fx[key + "_count"], fx[key + "_rgb"]), got)


@pytest.mark.parametrize("n,k,r", [(1, 1, 0.5), (2, 5, 3.0), (63, 10, 0.8), (64, 64, 4.0), (1000, 50, 0.3), (4097, 225, 0.4),
                                   (100000, 225, 0.05), (100000, 512, 0.2), (300000, 50, 0.02), (50000, 1, 0.1)])
def test_cuda_gather_matches_oracle_on_synthetic_maps(gpu_ctx, oracle, n, k, r):
    ph = synthetic_photons(n, seed=n + k)
    q = surface_queries(600, seed=k)
    gpu_ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    gpu_ctx.build_maps()
    for filt in (0, 1, 2):
        check(oracle.gather(ph, q, k, r, filt, threads=os.cpu_count() or 8), gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt))


def test_ties_duplicates_and_empty_map(gpu_ctx, oracle):
    q = surface_queries(200, seed=3)
    # empty map: nothing found, zero radiance
    gpu_ctx.load_photons(gi_b200.GI_MAP_GLOBAL, synthetic_photons(0, 1))
    gpu_ctx.build_maps()
    idx, cnt, rgb = gpu_ctx.gather(gi_b200.GI_MAP_GLOBAL, q, 20, 1.0)
    assert (cnt == 0).all() and (idx == -1).all() and (rgb == 0).all()
    # every photon four times (exact distance ties everywhere) + a cluster of identical positions
    base = synthetic_photons(3000, 5)
    ph = np.concatenate([base, base, base, base])
    ph["position"][100:400] = ph["position"][100]
    gpu_ctx.load_photons(gi_b200.GI_MAP_GLOBAL, ph)
    gpu_ctx.build_maps()
    for k in (1, 7, 50, 300):
        check(oracle.gather(ph, q, k, 0.5, 0), gpu_ctx.gather(gi_b200.GI_MAP_GLOBAL, q, k, 0.5, 0))
    # query exactly on photons (d^2 = 0) and far outside
    q2 = q.copy(); q2["p"][:100] = ph["position"][:100]; q2["p"][100:] += 50.0
    check(oracle.gather(ph, q2, 30, 0.3, 0), gpu_ctx.gather(gi_b200.GI_MAP_GLOBAL, q2, 30, 0.3, 0))


def test_photon_roundtrip_and_errors(gpu_ctx, tmp_path):
    ph = synthetic_photons(5000, 8)
    gpu_ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    assert gpu_ctx.photon_count(gi_b200.GI_MAP_CAUSTIC) == 5000
    assert gpu_ctx.get_photons(gi_b200.GI_MAP_CAUSTIC).tobytes() == ph.tobytes()
    p = str(tmp_path / "m.gipm")
    gpu_ctx.dump_photons_file(gi_b200.GI_MAP_CAUSTIC, p)
    assert read_photons(p).tobytes() == ph.tobytes()
    with pytest.raises(gi_b200.GiError, match="not built"):
        gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, surface_queries(4, 1), 5, 0.1)
    gpu_ctx.build_maps()
    with pytest.raises(gi_b200.GiError, match="estimate size"):
        gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, surface_queries(4, 1), 513, 0.1)


def test_full_size_properties(gpu_ctx):
    """10^7 photons (BASELINE default caustic map size), k = 225: too big for the brute-force oracle, so check
    size-independent properties: counts, sortedness, and that every reported photon lies within r and no unreported
    photon of a random sample is closer than the reported k-th."""
    n = 10_000_000
    ph = synthetic_photons(n, seed=77)
    gpu_ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    gpu_ctx.build_maps()
    q = surface_queries(2000, seed=12)
    k, r = 225, 0.02
    idx, cnt, rgb = gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, 0)
    assert (cnt == k).sum() > 1500
    rng = np.random.default_rng(0)
    for qi in rng.choice(len(q), 40, replace=False):
        ids = idx[qi][:cnt[qi]]
        assert (np.diff(ids) > 0).all()
        d2 = ((ph["position"] - q["p"][qi]) ** 2)
        d2 = (d2[:, 0] + d2[:, 1]) + d2[:, 2]
        inside = np.nonzero(d2 <= r * r)[0]
        if len(inside) <= k:
            assert np.array_equal(np.sort(inside), ids)
        else:
            order = inside[np.lexsort((inside, d2[inside]))][:k]
            assert np.array_equal(np.sort(order), ids)


def check_rgb(ref, got, rtol=1e-9):
    rcnt, rrgb = ref[1], ref[2]; cnt, rgb = got[1], got[2]
    assert np.array_equal(rcnt, cnt)
    scale = np.abs(rrgb).max(axis=1, keepdims=True)
    assert (np.abs(rgb - rrgb) <= rtol * scale + 1e-300).all()


@pytest.mark.parametrize("n,k,r,nq", [(100000, 225, 0.05, 6000), (20000, 50, 0.3, 5000), (3000, 512, 2.0, 4500),
                                      # maps of <= 8192 photons with k <= 102: the small-map kernels (chunk pools, a lane per query)
                                      (2176, 50, 2.5, 6000), (8000, 20, 0.12, 5000), (2000, 100, 0.45, 4200), (500, 50, 0.3, 4100)])
def test_radiance_only_path_on_sorted_batches(gpu_ctx, oracle, n, k, r, nq):
    """What the renderer and the bench run: no index output (the last cut marks instead of compacting, the candidate
    buffer is not a power of two) and batches of >= 4096 queries, which are walked along a Morton curve.  The small-map
    cases cover: every query full (r 2.5), many queries with fewer than k photons in range (r 0.12 / 0.3: the chunk
    centre's search comes back short), sparse queries whose shells overflow the per-lane lists (histogram path)."""
    ph = synthetic_photons(n, seed=7 * n + k)
    q = surface_queries(nq, seed=k + 1)
    gpu_ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    gpu_ctx.build_maps()
    for filt in (0, 1, 2):
        ref = oracle.gather(ph, q, k, r, filt, threads=os.cpu_count() or 8)
        check_rgb(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt, want_idx=False))
        check(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt))


def test_dense_blob_seen_from_a_distance(gpu_ctx, oracle):
    """A caustic map is dense in a few places.  For a query at some distance from such a blob every candidate lies in a
    thin shell of d^2: the cut's histogram must refine the crowded bin (levels) instead of falling back to ranking."""
    rng = np.random.default_rng(11)
    ph = synthetic_photons(62000, 21)
    ph["position"][2000:] = np.array([0.3, -1.0, 0.2]) + rng.normal(scale=[2e-3, 0.0, 2e-3], size=(60000, 3))   # blob on the floor
    q = surface_queries(512, seed=5)
    ang = rng.uniform(0, 2 * np.pi, 512); dist = rng.uniform(0.0, 0.25, 512)
    q["p"] = np.array([0.3, -1.0, 0.2]) + np.stack([np.cos(ang) * dist, np.zeros(512), np.sin(ang) * dist], axis=1)
    gpu_ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
    gpu_ctx.build_maps()
    for k, r in ((225, 0.225), (50, 0.4)):
        ref = oracle.gather(ph, q, k, r, 0, threads=os.cpu_count() or 8)
        check(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, 0))
        check_rgb(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, 0, want_idx=False))


# ---------------------------------------------------------------------------------------------- BASELINE-size maps
def _scene_queries(gpu_ctx, scene_gisc, n_side, ph, n_on_photons, seed):
    """gather queries as RayTrace sets them up: at the primary hits of an n_side x n_side camera grid, plus queries
    sitting on randomly chosen photons of the map (inside the dense caustic spots, where all k are found)"""
    from gi_testlib import queries_from_hits
    hs = gi_b200.HostScene(scene_gisc)
    gpu_ctx.upload_scene(hs)
    cam = hs.view.camera
    i, j = np.meshgrid(np.arange(float(n_side)), np.arange(float(n_side)), indexing="ij")
    dx = (2.0 * (i - n_side // 2) / n_side).reshape(-1, 1); dy = (2.0 * (j - n_side // 2) / n_side).reshape(-1, 1)
    eye = np.array(cam.eye[:])
    d = np.array(cam.towards[:]) + np.array(cam.right[:]) * np.tan(cam.xfov) * dx + np.array(cam.up[:]) * np.tan(cam.yfov) * dy
    d /= np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])[:, None]
    rays = np.concatenate([np.broadcast_to(eye, d.shape), d], axis=1)
    q = queries_from_hits(rays, gpu_ctx.intersect(rays))
    rng = np.random.default_rng(seed)
    extra = q[rng.integers(0, len(q), n_on_photons)].copy()
    extra["p"] = ph["position"][rng.integers(0, len(ph), n_on_photons)] + rng.normal(scale=1e-4, size=(n_on_photons, 3))
    return np.concatenate([q, extra])


def _reference_gather(tmp_path, gipm, q, k, r, filt):
    from gi_testlib import ref_harness, write_queries, read_ref_gather
    write_queries(tmp_path / "q.bin", q)
    ref_harness("gather", gipm, tmp_path / "q.bin", tmp_path / "g.bin", k, r, filt, os.cpu_count() or 8)
    cnt, rgb, idx = read_ref_gather(tmp_path / "g.bin")
    return idx, cnt, rgb


def _have_ref():
    from gi_testlib import have_ref_harness
    return have_ref_harness()


@pytest.mark.skipif(not _have_ref(), reason="needs oracle/_ref/ref_harness (built where the reference tree exists; travels to the GPU box)")
@pytest.mark.parametrize("case", ["cornell_1e7", "jensen_reftraced_3e6"])
def test_large_caustic_maps_against_live_reference_kdtree(gpu_ctx, tmp_path, case):
    """The kNN contract at the BASELINE sizes, against the reference's own R3Kdtree::FindClosestQuick + EstimateRadiance
    (R3Kdtree.cpp:688-848, photon_utils.cpp:72-162) run LIVE on the same map file: identical index sets and counts,
    radiance to 1e-9, k = 225 / 50, all three filters.
      cornell_1e7           cornell's DEFAULT caustic map (-caustic 10000000: ~1e7 photons, four rows of groups) traced by
                            gi_trace_photons (statistically the reference's map, tests/test_gpu_photons.py);
                            GI_SLOW_TESTS=1 lets the reference's MapPhotons trace it instead (minutes of CPU)
      jensen_reftraced_3e6  jensen.scn (configs[1]) with -caustic 3000000 traced by the reference's MapPhotons itself"""
    from gi_testlib import golden_gisc, ref_harness
    scene = case.split("_")[0]
    gipm = str(tmp_path / "caustic.gipm")
    if case == "jensen_reftraced_3e6" or os.environ.get("GI_SLOW_TESTS"):
        want = 3_000_000 if scene == "jensen" else 10_000_000
        ref_harness("photons", os.path.join(GOLDEN, "scn", scene + ".scn"), tmp_path / "ref", "-global", 2176, "-caustic", want,
                    "-threads", os.cpu_count() or 8)
        gipm = str(tmp_path / "ref.caustic.gipm")
        gpu_ctx.load_photons_file(gi_b200.GI_MAP_CAUSTIC, gipm)
        gpu_ctx.build_maps()
    else:
        want = 10_000_000
        gpu_ctx.upload_scene(gi_b200.HostScene(golden_gisc(scene)))
        p, _ = gi_b200.parse_flags(["-seed", "5", "-no_indirect"])
        st = gpu_ctx.trace_photons(p)
        assert st.caustic_stored >= want
        gpu_ctx.dump_photons_file(gi_b200.GI_MAP_CAUSTIC, gipm)
    ph = read_photons(gipm)
    assert len(ph) >= want
    q = _scene_queries(gpu_ctx, golden_gisc(scene), 40, ph, 600, seed=1)
    assert len(q) >= 2000
    full = 0
    for k, r, filt in ((225, 0.225, 0), (225, 0.225, 2), (50, 0.05, 1)):
        ref = _reference_gather(tmp_path, gipm, q, k, r, filt)
        check(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt))
        check_rgb(ref, gpu_ctx.gather(gi_b200.GI_MAP_CAUSTIC, q, k, r, filt, want_idx=False))
        full += int((ref[1] == k).sum())
    assert full > 1000                                  # the dense spots were really exercised


@pytest.mark.parametrize("mp,k,filt", [("caustic", 225, 0), ("caustic", 225, 1), ("caustic", 225, 2), ("caustic", 50, 0),
                                       ("global", 50, 0), ("global", 50, 2)])
def test_jensen_reference_maps(gpu_ctx, mp, k, filt):
    """jensen.scn (BASELINE configs[1]: rect light, glass + mirror sphere) maps traced by the reference's MapPhotons,
    gathers by the reference's kd-tree and estimator (tests/golden/jensen.gather.npz, make_jensen_golden.py)"""
    fx = np.load(os.path.join(GOLDEN, "jensen.gather.npz"))
    which = gi_b200.GI_MAP_CAUSTIC if mp == "caustic" else gi_b200.GI_MAP_GLOBAL
    gpu_ctx.load_photons(which, fx[mp + "_photons"].view(gi_b200.PHOTON).reshape(-1))
    gpu_ctx.build_maps()
    key = f"{mp}_{k}_{filt}"
    got = gpu_ctx.gather(which, fx["queries"].view(GI_QUERY).reshape(-1), k, float(fx[key + "_r"][0]), filt)
    check((fx[f"{mp}_{k}_idx"], fx[key + "_count"], fx[key + "_rgb"]), got)
