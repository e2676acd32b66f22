"""GPU suite (pytest -m gpu): gi_intersect_batch through the C ABI against (1) hit records of the reference itself
(golden fixtures), (2) the oracle restatement on larger seeded batches, (3) the reference binary live on a generated
large mesh.  Bar (BASELINE.json): bit-exact primitive / shape ids; we additionally require t, point and normal to be
bit-exact (the kernel keeps the reference's FP64 operation order), so the stated 1e-5 relative bound on t is met
with zero error."""
import glob
import os
import subprocess

import numpy as np
import pytest

import gi_b200
from gi_testlib import ROOT, REF_HARNESS, have_ref_harness, flatten_scene, random_rays, write_rays, read_ref_hits, golden_gisc

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = sorted(os.path.basename(p)[:-10] for p in glob.glob(os.path.join(GOLDEN, "*.isect.npz")))


def assert_same_hits(ref, got, what, check_prim=False):
    bad = np.nonzero(ref["hit"] != got["hit"])[0]
    assert len(bad) == 0, f"{what}: hit flags differ at rays {bad[:5]}"
    m = ref["hit"] == 1
    assert np.array_equal(ref["shape"][m], got["shape"][m]), f"{what}: shape ids differ"
    if check_prim:
        assert np.array_equal(ref["prim"][m], got["prim"][m]), f"{what}: primitive ids differ"
    t_rel = np.abs(ref["t"][m] - got["t"][m]) / np.maximum(np.abs(ref["t"][m]), 1e-300)
    assert (t_rel <= 1e-5).all()                       # the contract
    for f in ("t", "p", "n"):                          # what we actually achieve
        assert np.array_equal(ref[f][m], got[f][m]), f"{what}: {f} not bit-exact"


@pytest.mark.parametrize("name", SCENES)
def test_cuda_matches_reference_fixture(gpu_ctx, name):
    d = np.load(os.path.join(GOLDEN, name + ".isect.npz"))
    gpu_ctx.upload_scene(gi_b200.HostScene(golden_gisc(name)))
    got = gpu_ctx.intersect(d["rays"])
    assert_same_hits(d["hits"], got, name)


# violin: BASELINE.json configs[3], the one mesh large enough to stress the LBVH (104 300 triangles, deep Morton ties in
# the strings); the oracle (brute force like the reference) is pinned to the live reference on it by
# tests/test_oracle_isect.py::test_oracle_matches_live_reference_on_violin
@pytest.mark.parametrize("name", ["cornell", "jensen", "teapot", "stilllife", "stack", "art3", "art2", "circle", "violin"])
def test_cuda_matches_oracle_on_random_batches(gpu_ctx, oracle, name):
    gisc = golden_gisc(name)
    scene = gi_b200.HostScene(gisc)
    gpu_ctx.upload_scene(scene)
    lo = np.array(scene.view.bbox_min[:]); hi = np.array(scene.view.bbox_max[:])
    rays = random_rays(200000 if name != "teapot" else 40000, lo, hi, seed=21)
    if name == "violin":                                 # + camera rays: coherent, and they all aim at the instrument
        cam = scene.view.camera
        i, j = np.meshgrid(np.arange(160.0), np.arange(160.0), indexing="ij")
        dx = (2.0 * (i - 80) / 160).reshape(-1, 1); dy = (2.0 * (j - 80) / 160).reshape(-1, 1)
        eye = np.array(cam.eye[:])
        d = (np.array(cam.towards[:]) + np.array(cam.right[:]) * np.tan(cam.xfov) * dx + np.array(cam.up[:]) * np.tan(cam.yfov) * dy)
        d /= np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])[:, None]
        rays = np.concatenate([rays, np.concatenate([np.broadcast_to(eye, d.shape), d], axis=1)])
    h = oracle.load_scene(gisc)
    ref = oracle.intersect(h, rays, threads=os.cpu_count() or 8)
    got = gpu_ctx.intersect(rays)
    assert_same_hits(ref, got, name, check_prim=True)
    # secondary rays leaving the hit points (origins on / within 1e-6 of surfaces): negative-t rules, mesh poisoning
    m = ref["hit"] == 1
    rng = np.random.default_rng(4)
    d2 = rng.normal(size=(m.sum(), 3)); d2 /= np.sqrt((d2 * d2).sum(1))[:, None]
    d2 = np.where(((d2 * ref["n"][m]).sum(1) < 0)[:, None], -d2, d2)
    o2 = ref["p"][m] + 1e-6 * d2 * rng.choice([0.0, 1.0, 0.5], size=(m.sum(), 1))
    sec = np.concatenate([o2, d2], axis=1)
    ref2 = oracle.intersect(h, sec, threads=os.cpu_count() or 8)
    oracle.free_scene(h)
    assert_same_hits(ref2, gpu_ctx.intersect(sec), name + " secondary", check_prim=True)


def test_empty_and_degenerate_batches(gpu_ctx, oracle):
    gisc = os.path.join(GOLDEN, "cornell.gisc")
    gpu_ctx.upload_scene(gi_b200.HostScene(gisc))
    assert len(gpu_ctx.intersect(np.zeros((0, 6)))) == 0
    # a ray pointing away from everything, and axis-parallel rays (zero direction components, +-0)
    rays = np.array([[0, 0, 50, 0, 0, 1], [0, 0, 5, 0, 0, -1], [0.3, -0.2, 5, 0, 0, -1], [0.3, 0.2, 5, -0.0, 0, -1],
                     [0, 0.5, 0.5, 1, 0, 0], [0, 0.5, 0.5, 0, -1, 0], [0, 0.5, 0.5, 0, 1, -0.0]], dtype=np.float64)
    got = gpu_ctx.intersect(rays)
    assert got["hit"][0] == 0 and got["shape"][0] == -1 and got["t"][0] == 0
    assert got["hit"][1] == 1
    h = oracle.load_scene(gisc)
    assert_same_hits(oracle.intersect(h, rays), got, "axis-parallel", check_prim=True)
    oracle.free_scene(h)


def test_cone_scene_is_refused_loudly(gpu_ctx):
    """the reference's R3Intersects(ray, cone) is unfinished (zero normal, no cap, R3Isect.cpp:1201-1275): a scene
    holding a cone parses, but gi_upload_scene refuses it by name instead of rendering it as nothing"""
    hs = gi_b200.HostScene(os.path.join(GOLDEN, "scn", "cone.scn"))
    with pytest.raises(gi_b200.GiError, match="cone"):
        gpu_ctx.upload_scene(hs)
    gpu_ctx.upload_scene(gi_b200.HostScene(golden_gisc("cornell")))      # the context stays usable


def test_requires_a_scene():
    ctx = gi_b200.Context(0)
    with pytest.raises(gi_b200.GiError, match="no scene uploaded"):
        ctx.intersect(np.zeros((1, 6)))
    ctx.close()


def _write_torus(path_off, nu, nv):
    u = np.linspace(0, 2 * np.pi, nu, endpoint=False); v = np.linspace(0, 2 * np.pi, nv, endpoint=False)
    U, V = np.meshgrid(u, v, indexing="ij")
    R, r = 1.0, 0.35
    P = np.stack([(R + r * np.cos(V)) * np.cos(U), r * np.sin(V), (R + r * np.cos(V)) * np.sin(U)], axis=-1).reshape(-1, 3)
    with open(path_off, "w") as f:
        f.write(f"OFF\n{len(P)} {nu * nv} 0\n")
        for p in P:
            f.write("%.9g %.9g %.9g\n" % tuple(p))
        for i in range(nu):
            for j in range(nv):
                a = i * nv + j; b = ((i + 1) % nu) * nv + j; c = ((i + 1) % nu) * nv + (j + 1) % nv; d = i * nv + (j + 1) % nv
                f.write(f"4 {a} {b} {c} {d}\n")


@pytest.mark.skipif(not have_ref_harness(), reason="needs oracle/_ref/ref_harness (built where the reference tree exists)")
def test_large_mesh_against_live_reference(gpu_ctx, tmp_path):
    """a generated 2 x 160 x 100 = 32 000-triangle torus inside a scaled + rotated group, next to a sphere: the
    reference binary itself (brute force over every triangle) vs the BVH kernel, .scn/.off parsed by both sides"""
    _write_torus(tmp_path / "torus.off", 160, 100)
    (tmp_path / "torus.scn").write_text(
        "material 0 0 0  0.6 0.5 0.4  0.1 0.1 0.1  0 0 0  0 0 0  10 1 0\n"
        "begin 0  1.2 0 0.3 0.1   0 0.9 0 -0.2   -0.3 0 1.2 0   0 0 0 1\n  mesh 0 torus.off\nend\n"
        "sphere 0  0 0 0  0.4\n"
        "point_light 1 1 1  3 3 3  0 0 1\n")
    scn = str(tmp_path / "torus.scn"); gisc = str(tmp_path / "torus.gisc")
    flatten_scene(scn, gisc)
    scene = gi_b200.HostScene(scn)                      # the product's own .scn/.off loader
    gpu_ctx.upload_scene(scene)
    lo = np.array(scene.view.bbox_min[:]); hi = np.array(scene.view.bbox_max[:])
    rays = random_rays(3000, lo, hi, seed=9)
    write_rays(tmp_path / "rays.bin", rays)
    subprocess.run([REF_HARNESS, "isect", scn, str(tmp_path / "rays.bin"), str(tmp_path / "hits.bin"), str(os.cpu_count() or 8)],
                   check=True, capture_output=True)
    ref = read_ref_hits(tmp_path / "hits.bin")
    got = gpu_ctx.intersect(rays)
    assert ref["hit"].sum() > 500
    assert_same_hits(ref, got, "torus")
