"""CPU suite: the oracle restatement of R3Scene::Intersects (oracle/restate_isect.cpp) and the host simulation of
the device code (csrc/isect.cuh compiled for the CPU) against hit records produced by the REFERENCE ITSELF
(tests/golden/*.isect.npz, written by tests/golden/make_golden.py from oracle/_ref/ref_harness).
Bar: bit-exact -- hit flag, shape id, t, point and normal."""
import ctypes
import glob
import os

import numpy as np
import pytest

from gi_testlib import golden_gisc, GI_HIT, ROOT, REF_INPUT, have_ref_harness, ref_harness, flatten_scene, read_rays, write_rays, \
    read_ref_hits, random_rays

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = sorted(os.path.basename(p)[:-10] for p in glob.glob(os.path.join(GOLDEN, "*.isect.npz")))


def assert_same_hits(ref, got, what):
    assert np.array_equal(ref["hit"], got["hit"]), f"{what}: hit flags differ at {np.nonzero(ref['hit'] != got['hit'])[0][:5]}"
    m = ref["hit"] == 1
    assert np.array_equal(ref["shape"][m], got["shape"][m]), f"{what}: shape ids differ"
    for f in ("t", "p", "n"):
        assert np.array_equal(ref[f][m], got[f][m]), f"{what}: {f} not bit-exact"


@pytest.fixture(scope="module")
def sim():
    lib = ctypes.CDLL(os.path.join(ROOT, "tests", "host_sim", "libisect_sim.so"))
    lib.sim_scene_load.restype = ctypes.c_void_p
    lib.sim_scene_load.argtypes = [ctypes.c_char_p]
    lib.sim_scene_free.argtypes = [ctypes.c_void_p]
    lib.sim_intersect.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_int]
    return lib


def test_golden_set_is_complete():
    assert len(SCENES) >= 12 and "cornell" in SCENES and "teapot" in SCENES


@pytest.mark.parametrize("name", SCENES)
def test_oracle_matches_reference(oracle, name):
    d = np.load(os.path.join(GOLDEN, name + ".isect.npz"))
    h = oracle.load_scene(golden_gisc(name))
    got = oracle.intersect(h, d["rays"], threads=4)
    oracle.free_scene(h)
    assert_same_hits(d["hits"], got, name)


@pytest.mark.parametrize("name", SCENES)
def test_device_code_on_cpu_matches_reference(sim, name):
    d = np.load(os.path.join(GOLDEN, name + ".isect.npz"))
    s = sim.sim_scene_load(golden_gisc(name).encode())
    assert s
    rays = np.ascontiguousarray(d["rays"], dtype=np.float64)
    got = np.zeros(len(rays), dtype=GI_HIT)
    assert sim.sim_intersect(s, rays.ctypes.data, len(rays), got.ctypes.data, 4) == 0
    sim.sim_scene_free(s)
    assert_same_hits(d["hits"], got, name)


@pytest.mark.skipif(not (have_ref_harness() and os.path.isdir(REF_INPUT)), reason="needs the reference tree")
def test_oracle_matches_live_reference_on_violin(oracle, tmp_path):
    """the one big mesh (104 k triangles) has no committed fixture: compare live, few rays (the reference is brute force)"""
    scn = os.path.join(REF_INPUT, "violin.scn")
    gisc = str(tmp_path / "violin.gisc")
    flatten_scene(scn, gisc)
    ref_harness("camrays", scn, 12, 12, tmp_path / "cam.bin")
    rays = read_rays(tmp_path / "cam.bin")
    tail = np.frombuffer(open(gisc, "rb").read()[-128:], dtype="<f8")
    rays = np.concatenate([rays, random_rays(150, tail[6:9], tail[9:12], seed=2)])
    write_rays(tmp_path / "rays.bin", rays)
    ref_harness("isect", scn, tmp_path / "rays.bin", tmp_path / "hits.bin", 8)
    ref = read_ref_hits(tmp_path / "hits.bin")
    h = oracle.load_scene(gisc)
    got = oracle.intersect(h, rays, threads=8)
    oracle.free_scene(h)
    assert_same_hits(ref, got, "violin")
