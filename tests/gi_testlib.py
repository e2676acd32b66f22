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
GOLDEN = os.path.join(ROOT, "tests", "golden")

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


def golden_gisc(name):
    """path of the flattened scene tests/golden/<name>.gisc; large ones are committed gzip'd and unpacked on first use
    (the unpacked copy is git-ignored)"""
    p = os.path.join(GOLDEN, name + ".gisc")
    if not os.path.exists(p) and os.path.exists(p + ".gz"):
        import gzip
        tmp = p + ".tmp%d" % os.getpid()
        with gzip.open(p + ".gz", "rb") as f, open(tmp, "wb") as g:
            g.write(f.read())
        os.replace(tmp, p)
    return p


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


# ---------------------------------------------------------------------------------------------- photon maps
GI_PHOTON = np.dtype([("position", "<f8", 3), ("rgbe", "u1", 4), ("direction", "<u2"), ("pad", "<u2")])
GI_QUERY = np.dtype([("p", "<f8", 3), ("n", "<f8", 3), ("bounce", "<f8", 3), ("cos_theta", "<f8"), ("kd", "<f8", 3),
                     ("ks", "<f8", 3), ("shininess", "<f8")])
assert GI_PHOTON.itemsize == 32 and GI_QUERY.itemsize == 136
PMAP_MAGIC = 0x4D504947
QRY_MAGIC = 0x59525147


def read_photons(path):
    with open(path, "rb") as f:
        hdr = np.frombuffer(f.read(8), dtype="<u4")
        assert hdr[0] == PMAP_MAGIC
        n = int(np.frombuffer(f.read(8), dtype="<u8")[0])
        return np.frombuffer(f.read(32 * n), dtype=GI_PHOTON).copy()


def write_photons(path, ph):
    ph = np.ascontiguousarray(ph, dtype=GI_PHOTON)
    with open(path, "wb") as f:
        f.write(np.array([PMAP_MAGIC, 1], dtype="<u4").tobytes())
        f.write(np.array([len(ph)], dtype="<u8").tobytes())
        f.write(ph.tobytes())


def write_queries(path, q):
    q = np.ascontiguousarray(q, dtype=GI_QUERY)
    with open(path, "wb") as f:
        f.write(np.array([QRY_MAGIC, len(q)], dtype="<u4").tobytes())
        f.write(q.tobytes())


def read_ref_gather(path):
    """output of `ref_harness gather`: (count[n], rgb[n,3], idx[n,k])"""
    with open(path, "rb") as f:
        hdr = np.frombuffer(f.read(16), dtype="<u4")
        assert hdr[0] == 0x54414747
        n, k = int(hdr[1]), int(hdr[2])
        rec = np.frombuffer(f.read(32 * n), dtype=np.dtype([("count", "<i4"), ("pad", "<i4"), ("rgb", "<f8", 3)]))
        idx = np.frombuffer(f.read(4 * n * k), dtype="<i4").reshape(n, k)
        return rec["count"].copy(), rec["rgb"].copy(), idx.copy()


def queries_from_hits(rays, hits, kd=(0.5, 0.6, 0.7), ks=(0.2, 0.1, 0.3), shininess=10.0):
    """gather queries at hit points, set up as RayTrace / CausticIllumination do (raytracer.cpp:138-149,174-190):
    view = ray direction, cos_theta = n . (-view), exact_bounce = mirror direction of the view ray"""
    m = hits["hit"] == 1
    d = rays[m, 3:6]; n = hits["n"][m]
    cos_theta = -(n * d).sum(1)
    q = np.zeros(m.sum(), dtype=GI_QUERY)
    q["p"] = hits["p"][m]; q["n"] = n
    q["bounce"] = d + 2.0 * cos_theta[:, None] * n
    q["cos_theta"] = cos_theta
    q["kd"] = kd; q["ks"] = ks; q["shininess"] = shininess
    return q


def synthetic_photons(n, seed, lo=(-1, -1, -1), hi=(1, 1, 1)):
    """photons uniform on the six faces of a box, random power / direction codes"""
    rng = np.random.default_rng(seed)
    lo = np.asarray(lo, dtype=np.float64); hi = np.asarray(hi, dtype=np.float64)
    p = lo + rng.random((n, 3)) * (hi - lo)
    face = rng.integers(0, 6, n)
    ax = face % 3
    p[np.arange(n), ax] = np.where(face < 3, lo[ax], hi[ax])
    ph = np.zeros(n, dtype=GI_PHOTON)
    ph["position"] = p
    ph["rgbe"][:, :3] = rng.integers(0, 256, (n, 3)); ph["rgbe"][:, 3] = rng.integers(110, 130, n)
    ph["direction"] = rng.integers(0, 65536, n)
    return ph


def _oracle_gather(self, photons, queries, k, r, filter=0, fa=0.918, fb=1.953, fk=1.0, threads=8, want_idx=True):
    photons = np.ascontiguousarray(photons, dtype=GI_PHOTON); queries = np.ascontiguousarray(queries, dtype=GI_QUERY)
    nq = len(queries)
    idx = np.full((nq, k), -1, dtype=np.int32) if want_idx else None
    cnt = np.zeros(nq, dtype=np.int32); rgb = np.zeros((nq, 3), dtype=np.float64)
    f = self.lib.oracle_gather
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_double,
                  ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_int]
    rc = f(photons.ctypes.data, len(photons), queries.ctypes.data, nq, k, r, filter, fa, fb, fk,
           idx.ctypes.data if want_idx else None, cnt.ctypes.data, rgb.ctypes.data, threads)
    assert rc == 0
    return idx, cnt, rgb


Oracle.gather = _oracle_gather


def _oracle_irradiance_prepass(self, photons, k, r, threads=8):
    """colours (n x 4 RGBE bytes) of the photons after the -cache pre-pass (oracle/restate_cache.cpp)"""
    photons = np.ascontiguousarray(photons, dtype=GI_PHOTON)
    out = np.zeros((len(photons), 4), dtype=np.uint8)
    f = self.lib.oracle_irradiance_prepass
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_int]
    assert f(photons.ctypes.data, len(photons), k, r, out.ctypes.data, threads) == 0
    return out


def _oracle_cached_radiance(self, photons, queries, r, threads=8):
    """EstimateCachedRadiance per query on an already converted map: (idx[n], rgb[n, 3])"""
    photons = np.ascontiguousarray(photons, dtype=GI_PHOTON); queries = np.ascontiguousarray(queries, dtype=GI_QUERY)
    idx = np.full(len(queries), -1, dtype=np.int32); rgb = np.zeros((len(queries), 3), dtype=np.float64)
    f = self.lib.oracle_cached_radiance
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_double, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_int]
    assert f(photons.ctypes.data, len(photons), queries.ctypes.data, len(queries), r, idx.ctypes.data, rgb.ctypes.data, threads) == 0
    return idx, rgb


Oracle.irradiance_prepass = _oracle_irradiance_prepass
Oracle.cached_radiance = _oracle_cached_radiance


def read_png_rgb(path):
    """Minimal PNG reader (8-bit RGB / RGBA, non-interlaced, all five filter types): returns uint8 [H, W, 3], top row first."""
    import struct
    import zlib
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n", "not a PNG"
    pos, idat, ihdr = 8, b"", None
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        if typ == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    w, h, depth, ctype, _, _, interlace = ihdr
    assert depth == 8 and ctype in (2, 6) and interlace == 0
    bpp = 3 if ctype == 2 else 4
    data = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + w * bpp)
    out = np.zeros((h, w * bpp), dtype=np.int32)
    prev = np.zeros(w * bpp, dtype=np.int32)
    for y in range(h):
        ft, line = int(data[y, 0]), data[y, 1:].astype(np.int32)
        cur = np.zeros(w * bpp, dtype=np.int32)
        if ft == 0:
            cur = line.copy()
        elif ft == 2:
            cur = (line + prev) & 255
        else:
            for x in range(w * bpp):
                a = cur[x - bpp] if x >= bpp else 0
                b = prev[x]
                c = prev[x - bpp] if x >= bpp else 0
                if ft == 1:
                    pred = a
                elif ft == 3:
                    pred = (a + b) >> 1
                else:
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[x] = (line[x] + pred) & 255
        out[y] = cur
        prev = cur
    return out.reshape(h, w, bpp)[:, :, :3].astype(np.uint8)
