"""Shared helpers for the tests and bench: file formats of oracle/ref_harness, ctypes bindings.

Test infrastructure only.  `Oracle` wraps oracle/liboracle.so (our CPU restatement) and `RefHarness`
wraps oracle/_ref/ref_harness (the unmodified reference sources compiled in place) -- both are
checkers; the product is global-illumination_b200/libgi_b200.so (`gi_b200` package).
"""
import ctypes
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "global-illumination_b200")
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
REF_INPUT = "/root/reference/input"
SCENE_TOOL = os.path.join(PKG, "bin", "gi_scene_tool")
SCENES_DIR = os.path.join(ROOT, "tests", "golden", "scenes")

RAYS_MAGIC = 0x59415247
HITS_MAGIC = 0x54494847

# ref_harness HitRec (64 bytes)
REF_HIT = np.dtype([("hit", "<i4"), ("shape", "<i4"), ("t", "<f8"), ("p", "<f8", 3), ("n", "<f8", 3)])
# GiHit of include/gi_b200.h (72 bytes)
GI_HIT = np.dtype([("hit", "<i4"), ("shape", "<i4"), ("prim", "<i4"), ("element", "<i4"), ("t", "<f8"),
                   ("p", "<f8", 3), ("n", "<f8", 3)])
assert REF_HIT.itemsize == 64 and GI_HIT.itemsize == 72


def write_rays(path, rays):
    rays = np.ascontiguousarray(rays, dtype="<f8").reshape(-1, 6)
    with open(path, "wb") as f:
        f.write(np.array([RAYS_MAGIC, len(rays)], dtype="<u4").tobytes())
        f.write(rays.tobytes())


def read_rays(path):
    with open(path, "rb") as f:
        hdr = np.frombuffer(f.read(8), dtype="<u4")
        assert hdr[0] == RAYS_MAGIC
        return np.frombuffer(f.read(), dtype="<f8").reshape(-1, 6).copy()


def read_ref_hits(path):
    with open(path, "rb") as f:
        hdr = np.frombuffer(f.read(8), dtype="<u4")
        assert hdr[0] == HITS_MAGIC
        return np.frombuffer(f.read(), dtype=REF_HIT).copy()


def have_ref_harness():
    return os.path.exists(REF_HARNESS)


def ref_harness(*args, check=True):
    return subprocess.run([REF_HARNESS, *map(str, args)], check=check, capture_output=True, text=True)


def flatten_scene(scn_path, gisc_path, real=False):
    args = [SCENE_TOOL, "flatten", scn_path, gisc_path] + (["-real"] if real else [])
    subprocess.run(args, check=True)


def random_rays(n, lo, hi, seed, scale=1.5):
    """origins uniform in `scale` x the scene bbox, directions uniform on the sphere."""
    rng = np.random.default_rng(seed)
    lo = np.asarray(lo, dtype=np.float64); hi = np.asarray(hi, dtype=np.float64)
    c = 0.5 * (lo + hi); h = 0.5 * (hi - lo) * scale
    o = c + (rng.random((n, 3)) * 2 - 1) * h
    d = rng.normal(size=(n, 3))
    d /= np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]))[:, None]
    return np.concatenate([o, d], axis=1)


class Oracle:
    """ctypes view of oracle/liboracle.so (CPU restatement)."""

    def __init__(self):
        self.lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
        self.lib.oracle_scene_load.restype = ctypes.c_void_p
        self.lib.oracle_scene_load.argtypes = [ctypes.c_char_p]
        self.lib.oracle_scene_free.argtypes = [ctypes.c_void_p]
        self.lib.oracle_intersect.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p,
                                              ctypes.c_int]

    def load_scene(self, gisc_path):
        h = self.lib.oracle_scene_load(gisc_path.encode())
        if not h:
            raise RuntimeError("oracle: cannot load " + gisc_path)
        return h

    def free_scene(self, h):
        self.lib.oracle_scene_free(h)

    def intersect(self, scene, rays, threads=8):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        hits = np.zeros(len(rays), dtype=GI_HIT)
        rc = self.lib.oracle_intersect(scene, rays.ctypes.data, len(rays), hits.ctypes.data, threads)
        assert rc == 0
        return hits
