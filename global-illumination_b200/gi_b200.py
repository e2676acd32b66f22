"""ctypes binding of libgi_b200.so (include/gi_b200.h) -- the thin Python host side used by tests and bench.py.

The product is the C-ABI library; this module only marshals numpy buffers into it.  There is NO CPU path: if the
shared library is missing, or no CUDA device is present, everything here raises.
Structure layouts mirror include/gi_b200.h one to one.
"""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgi_b200.so")

GI_MAP_GLOBAL, GI_MAP_CAUSTIC = 0, 1
GI_FILTER_DISK, GI_FILTER_CONE, GI_FILTER_GAUSS = 0, 1, 2
GI_VIEW_GLOBAL_PHOTONS, GI_VIEW_CAUSTIC_PHOTONS, GI_VIEW_RAYS, GI_VIEW_PATHS = 0, 1, 2, 3
NOALLOC_MAX_QUERIES = 262144      # GI_NOALLOC_MAX_QUERIES of include/gi_b200.h

RAY = np.dtype([("o", "<f8", 3), ("d", "<f8", 3)])
MATERIAL = np.dtype([("ka", "<f8", 3), ("kd", "<f8", 3), ("ks", "<f8", 3), ("kt", "<f8", 3), ("ke", "<f8", 3), ("shininess", "<f8"), ("ir", "<f8")])
ELEMENT = np.dtype([("node", "<i4"), ("material", "<i4"), ("bbox", "<f8", 6)])
HIT_COMPACT = np.dtype([("prim", "<i4"), ("shape", "<i4"), ("t", "<f8")])
QUERY_COMPACT = np.dtype([("p", "<f8", 3), ("n", "<f8", 3), ("view", "<f8", 3), ("material", "<i4"), ("pad", "<i4")])
HIT = np.dtype([("hit", "<i4"), ("shape", "<i4"), ("prim", "<i4"), ("element", "<i4"), ("t", "<f8"),
                ("p", "<f8", 3), ("n", "<f8", 3)])
PHOTON = np.dtype([("position", "<f8", 3), ("rgbe", "u1", 4), ("direction", "<u2"), ("pad", "<u2")])
QUERY = np.dtype([("p", "<f8", 3), ("n", "<f8", 3), ("bounce", "<f8", 3), ("cos_theta", "<f8"), ("kd", "<f8", 3),
                  ("ks", "<f8", 3), ("shininess", "<f8")])


class GiCamera(ctypes.Structure):
    _fields_ = [("eye", ctypes.c_double * 3), ("towards", ctypes.c_double * 3), ("up", ctypes.c_double * 3),
                ("right", ctypes.c_double * 3), ("xfov", ctypes.c_double), ("yfov", ctypes.c_double),
                ("neardist", ctypes.c_double), ("fardist", ctypes.c_double)]


class GiScene(ctypes.Structure):
    _fields_ = [("n_nodes", ctypes.c_int32), ("n_elements", ctypes.c_int32), ("n_materials", ctypes.c_int32),
                ("n_prims", ctypes.c_int32), ("n_lights", ctypes.c_int32), ("n_shapes", ctypes.c_int32),
                ("nodes", ctypes.c_void_p), ("elements", ctypes.c_void_p), ("materials", ctypes.c_void_p),
                ("prims", ctypes.c_void_p), ("shapes", ctypes.c_void_p), ("lights", ctypes.c_void_p),
                ("camera", GiCamera), ("ambient", ctypes.c_double * 3), ("background", ctypes.c_double * 3),
                ("bbox_min", ctypes.c_double * 3), ("bbox_max", ctypes.c_double * 3), ("centroid", ctypes.c_double * 3),
                ("radius", ctypes.c_double)]


class GiParams(ctypes.Structure):
    _fields_ = [("verbose", ctypes.c_int32), ("threads", ctypes.c_int32), ("fresnel", ctypes.c_int32),
                ("ir_air", ctypes.c_double),
                ("ambient", ctypes.c_int32), ("direct_illum", ctypes.c_int32), ("transmissive_illum", ctypes.c_int32),
                ("specular_illum", ctypes.c_int32), ("indirect_illum", ctypes.c_int32), ("caustic_illum", ctypes.c_int32),
                ("direct_photon_illum", ctypes.c_int32), ("fast_global", ctypes.c_int32), ("irradiance_cache", ctypes.c_int32),
                ("shadows", ctypes.c_int32), ("soft_shadows", ctypes.c_int32), ("light_test", ctypes.c_int32),
                ("shadow_test", ctypes.c_int32),
                ("monte_carlo", ctypes.c_int32), ("max_monte_depth", ctypes.c_int32),
                ("prob_absorb", ctypes.c_double),
                ("recursive_shadows", ctypes.c_int32), ("distrib_transmissive", ctypes.c_int32),
                ("transmissive_test", ctypes.c_int32), ("distrib_specular", ctypes.c_int32), ("specular_test", ctypes.c_int32),
                ("depth_of_field", ctypes.c_int32), ("dof_test", ctypes.c_int32),
                ("focus_depth", ctypes.c_double), ("aperture_radius", ctypes.c_double),
                ("global_photon_count", ctypes.c_int32), ("caustic_photon_count", ctypes.c_int32),
                ("max_photon_depth", ctypes.c_int32),
                ("indirect_test", ctypes.c_int32), ("global_estimate_size", ctypes.c_int32),
                ("global_estimate_dist", ctypes.c_double),
                ("global_filter", ctypes.c_int32), ("caustic_estimate_size", ctypes.c_int32),
                ("caustic_estimate_dist", ctypes.c_double),
                ("caustic_filter", ctypes.c_int32),
                ("filter_const_a", ctypes.c_double), ("filter_const_b", ctypes.c_double), ("filter_const_k", ctypes.c_double),
                ("seed", ctypes.c_uint64), ("rank", ctypes.c_int32), ("world_size", ctypes.c_int32)]


class GiPhotonStats(ctypes.Structure):
    _fields_ = [("global_stored", ctypes.c_uint64), ("caustic_stored", ctypes.c_uint64),
                ("global_emitted", ctypes.c_uint64), ("caustic_emitted", ctypes.c_uint64),
                ("indirect_illum", ctypes.c_int32), ("caustic_illum", ctypes.c_int32),
                ("direct_photon_illum", ctypes.c_int32), ("rounds", ctypes.c_int32),
                ("trace_ms", ctypes.c_double), ("build_ms", ctypes.c_double),
                ("total_power", ctypes.c_double), ("cache_ms", ctypes.c_double)]


class GiRenderStats(ctypes.Structure):
    _fields_ = [("screen_rays", ctypes.c_uint64), ("shadow_rays", ctypes.c_uint64), ("monte_rays", ctypes.c_uint64),
                ("transmissive_samples", ctypes.c_uint64), ("specular_samples", ctypes.c_uint64),
                ("indirect_samples", ctypes.c_uint64), ("caustic_samples", ctypes.c_uint64),
                ("render_ms", ctypes.c_double)]


class GiStageTimes(ctypes.Structure):
    _fields_ = [(n, ctypes.c_double) for n in
                ("emit_filter_ms", "trace_ms", "photon_exchange_ms", "photon_sort_ms", "kd_build_ms",
                 "primary_ms", "direct_ms", "path_begin_ms", "path_hit_ms", "path_shade_ms", "path_finish_ms", "gather_global_ms",
                 "gather_caustic_ms", "commit_ms", "resolve_ms", "frame_exchange_ms")]


# every symbol include/gi_b200.h declares (tests check the built library exports all of them)
EXPORTS = ["gi_abi_version", "gi_last_error", "gi_create", "gi_destroy", "gi_default_params", "gi_launch_count",
           "gi_queue_overflow_count", "gi_cached_radiance_batch", "gi_cached_radiance_batch_device",
           "gi_host_scene_load", "gi_host_scene_view", "gi_host_scene_save", "gi_host_scene_free",
           "gi_upload_scene", "gi_intersect_batch", "gi_intersect_batch_device",
           "gi_load_photons", "gi_load_photons_device", "gi_load_photons_file", "gi_dump_photons_file",
           "gi_photon_count", "gi_get_photons", "gi_get_photons_device", "gi_build_maps", "gi_gather_batch", "gi_gather_batch_device",
           "gi_gather_batch_device_noalloc", "gi_stage_timing", "gi_stage_times", "gi_load_photons_sharded_device", "gi_debug_view", "gi_intersect_batch_compact", "gi_gather_batch_compact",
           "gi_trace_photons", "gi_scale_photon_power", "gi_render", "gi_render_float",
           "gi_default_cli_options", "gi_parse_args", "gi_write_image",
           "gi_comm_unique_id", "gi_comm_init_rank", "gi_comm_init_all", "gi_comm_info", "gi_comm_destroy"]

_lib = None


def lib():
    """Load libgi_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libgi_b200.so is not built (run `make lib` or __graft_entry__.build()); "
                               "gi_b200 has no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        L.gi_last_error.restype = ctypes.c_char_p
        vp, u64, i32, dbl = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_double
        L.gi_create.argtypes = [i32, ctypes.POINTER(vp)]
        L.gi_destroy.argtypes = [vp]; L.gi_destroy.restype = None
        L.gi_default_params.argtypes = [ctypes.POINTER(GiParams)]; L.gi_default_params.restype = None
        L.gi_launch_count.argtypes = [vp, ctypes.POINTER(u64)]
        L.gi_stage_timing.argtypes = [vp, i32]
        L.gi_intersect_batch_compact.argtypes = [vp, vp, u64, vp]
        L.gi_gather_batch_compact.argtypes = [vp, i32, vp, u64, i32, dbl, i32, ctypes.POINTER(GiParams), vp, vp]
        L.gi_debug_view.argtypes = [vp, i32, ctypes.POINTER(GiParams), i32, i32, i32, vp]
        L.gi_stage_times.argtypes = [vp, ctypes.POINTER(GiStageTimes)]
        L.gi_queue_overflow_count.argtypes = [vp, ctypes.POINTER(u64)]
        L.gi_host_scene_load.argtypes = [ctypes.c_char_p, i32, ctypes.POINTER(vp)]
        L.gi_host_scene_view.argtypes = [vp, ctypes.POINTER(GiScene)]
        L.gi_host_scene_save.argtypes = [vp, ctypes.c_char_p]
        L.gi_host_scene_free.argtypes = [vp]; L.gi_host_scene_free.restype = None
        L.gi_upload_scene.argtypes = [vp, ctypes.POINTER(GiScene)]
        L.gi_intersect_batch.argtypes = [vp, vp, u64, vp]
        L.gi_intersect_batch_device.argtypes = [vp, vp, u64, vp, vp]
        L.gi_load_photons.argtypes = [vp, i32, vp, u64]
        L.gi_load_photons_device.argtypes = [vp, i32, vp, u64]
        L.gi_load_photons_sharded_device.argtypes = [vp, i32, vp, u64]
        L.gi_load_photons_file.argtypes = [vp, i32, ctypes.c_char_p]
        L.gi_dump_photons_file.argtypes = [vp, i32, ctypes.c_char_p]
        L.gi_photon_count.argtypes = [vp, i32, ctypes.POINTER(u64)]
        L.gi_get_photons.argtypes = [vp, i32, vp, u64]
        L.gi_get_photons_device.argtypes = [vp, i32, vp, u64, vp]
        L.gi_build_maps.argtypes = [vp, ctypes.POINTER(GiParams)]
        L.gi_gather_batch.argtypes = [vp, i32, vp, u64, i32, dbl, i32, ctypes.POINTER(GiParams), vp, vp, vp]
        L.gi_gather_batch_device.argtypes = [vp, i32, vp, u64, i32, dbl, i32, ctypes.POINTER(GiParams), vp, vp, vp, vp]
        L.gi_gather_batch_device_noalloc.argtypes = [vp, i32, vp, u64, i32, dbl, i32, ctypes.POINTER(GiParams), vp, vp, vp, vp]
        L.gi_cached_radiance_batch.argtypes = [vp, vp, u64, dbl, vp]
        L.gi_cached_radiance_batch_device.argtypes = [vp, vp, u64, dbl, vp, vp]
        L.gi_trace_photons.argtypes = [vp, ctypes.POINTER(GiParams), ctypes.POINTER(GiPhotonStats)]
        L.gi_scale_photon_power.argtypes = [vp, i32, dbl]
        L.gi_render.argtypes = [vp, ctypes.POINTER(GiParams), i32, i32, i32, vp, ctypes.POINTER(GiRenderStats)]
        L.gi_render_float.argtypes = [vp, ctypes.POINTER(GiParams), i32, i32, i32, vp, ctypes.POINTER(GiRenderStats)]
        L.gi_comm_unique_id.argtypes = [vp]
        L.gi_comm_init_rank.argtypes = [vp, vp, i32, i32]
        L.gi_comm_init_all.argtypes = [ctypes.POINTER(vp), i32]
        L.gi_comm_info.argtypes = [vp, ctypes.POINTER(i32), ctypes.POINTER(i32)]
        L.gi_comm_destroy.argtypes = [vp]
        _lib = L
    return _lib


class GiError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise GiError(lib().gi_last_error().decode(errors="replace"))


def comm_unique_id():
    """128 bytes (an ncclUniqueId) made by rank 0; hand them to every rank's Context.comm_init_rank"""
    buf = ctypes.create_string_buffer(128)
    _check(lib().gi_comm_unique_id(buf))
    return buf.raw


def comm_init_all(contexts):
    """one process driving several GPUs: rank i = contexts[i] (ncclCommInitAll)"""
    arr = (ctypes.c_void_p * len(contexts))(*[c.handle for c in contexts])
    _check(lib().gi_comm_init_all(arr, len(contexts)))


def default_params():
    p = GiParams()
    lib().gi_default_params(ctypes.byref(p))
    return p


class HostScene:
    """Flattened scene on the host: the .scn/.off loader of the reference's CLI, or a saved .gisc file."""

    def __init__(self, path, real_material=False):
        self.handle = ctypes.c_void_p()
        _check(lib().gi_host_scene_load(path.encode(), int(real_material), ctypes.byref(self.handle)))
        self.view = GiScene()
        _check(lib().gi_host_scene_view(self.handle, ctypes.byref(self.view)))

    def save(self, gisc_path):
        _check(lib().gi_host_scene_save(self.handle, gisc_path.encode()))

    def close(self):
        if self.handle:
            lib().gi_host_scene_free(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def materials(self):
        """GiScene.materials as a numpy record array (MATERIAL)"""
        n = self.view.n_materials
        return np.frombuffer((ctypes.c_char * (n * MATERIAL.itemsize)).from_address(self.view.materials), dtype=MATERIAL).copy()

    def elements(self):
        """GiScene.elements as a numpy record array (ELEMENT): node, material index, bbox"""
        n = self.view.n_elements
        return np.frombuffer((ctypes.c_char * (n * ELEMENT.itemsize)).from_address(self.view.elements), dtype=ELEMENT).copy()


class Context:
    """One GPU.  Mirrors the gi_* entry points; host numpy buffers in, host numpy buffers out."""

    def __init__(self, device=0):
        self.handle = ctypes.c_void_p()
        _check(lib().gi_create(device, ctypes.byref(self.handle)))

    def close(self):
        if self.handle:
            lib().gi_destroy(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def comm_init_rank(self, unique_id, rank, world):
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        _check(lib().gi_comm_init_rank(self.handle, buf, rank, world))

    def comm_info(self):
        r, w = ctypes.c_int(), ctypes.c_int()
        _check(lib().gi_comm_info(self.handle, ctypes.byref(r), ctypes.byref(w)))
        return r.value, w.value

    def comm_destroy(self):
        _check(lib().gi_comm_destroy(self.handle))

    def launch_count(self):
        n = ctypes.c_uint64()
        _check(lib().gi_launch_count(self.handle, ctypes.byref(n)))
        return n.value

    def stage_timing(self, on=True):
        _check(lib().gi_stage_timing(self.handle, 1 if on else 0))

    def stage_times(self):
        """per-stage device milliseconds of the last trace_photons / render of this context (dict)"""
        t = GiStageTimes()
        _check(lib().gi_stage_times(self.handle, ctypes.byref(t)))
        return {n: float(getattr(t, n)) for n, _ in GiStageTimes._fields_}

    def debug_view(self, kind, width, height, draw_shapes=True, params=None):
        """one of the reference's debug views (visualize.cpp) as a uint8 array [H, W, 3], row 0 = bottom row"""
        out = np.zeros((height, width, 3), dtype=np.uint8)
        p = params or default_params()
        _check(lib().gi_debug_view(self.handle, kind, ctypes.byref(p), width, height, 1 if draw_shapes else 0, out.ctypes.data))
        return out

    def queue_overflow_count(self):
        n = ctypes.c_uint64()
        _check(lib().gi_queue_overflow_count(self.handle, ctypes.byref(n)))
        return n.value

    def upload_scene(self, scene):
        _check(lib().gi_upload_scene(self.handle, ctypes.byref(scene.view)))

    def intersect(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        hits = np.zeros(len(rays), dtype=HIT)
        _check(lib().gi_intersect_batch(self.handle, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def intersect_compact(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        hits = np.zeros(len(rays), dtype=HIT_COMPACT)
        _check(lib().gi_intersect_batch_compact(self.handle, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def gather_compact(self, which, queries, k, r, filter=GI_FILTER_DISK, params=None):
        """queries: QUERY_COMPACT records (point, normal, view direction, material index of the uploaded scene)"""
        queries = np.ascontiguousarray(queries, dtype=QUERY_COMPACT)
        n = len(queries)
        cnt = np.zeros(n, dtype=np.int32); rgb = np.zeros((n, 3), dtype=np.float64)
        p = params or default_params()
        _check(lib().gi_gather_batch_compact(self.handle, which, queries.ctypes.data, n, k, float(r), filter, ctypes.byref(p),
                                             cnt.ctypes.data, rgb.ctypes.data))
        return cnt, rgb

    def intersect_device(self, d_rays, n, d_hits, stream=None):
        _check(lib().gi_intersect_batch_device(self.handle, d_rays, n, d_hits, stream))

    def load_photons(self, which, photons):
        photons = np.ascontiguousarray(photons, dtype=PHOTON)
        _check(lib().gi_load_photons(self.handle, which, photons.ctypes.data, len(photons)))

    def load_photons_device(self, which, d_ptr, n):
        _check(lib().gi_load_photons_device(self.handle, which, d_ptr, n))

    def load_photons_sharded_device(self, which, d_ptr, n_share):
        """collective: this rank's share (device pointer); the map holds the shares of all ranks in rank order"""
        _check(lib().gi_load_photons_sharded_device(self.handle, which, d_ptr, n_share))

    def load_photons_file(self, which, path):
        _check(lib().gi_load_photons_file(self.handle, which, path.encode()))

    def dump_photons_file(self, which, path):
        _check(lib().gi_dump_photons_file(self.handle, which, path.encode()))

    def photon_count(self, which):
        n = ctypes.c_uint64()
        _check(lib().gi_photon_count(self.handle, which, ctypes.byref(n)))
        return n.value

    def get_photons(self, which):
        out = np.zeros(self.photon_count(which), dtype=PHOTON)
        _check(lib().gi_get_photons(self.handle, which, out.ctypes.data, len(out)))
        return out

    def get_photons_device(self, which, d_out, capacity, stream=None):
        _check(lib().gi_get_photons_device(self.handle, which, d_out, capacity, stream))

    def build_maps(self, params=None):
        p = params or default_params()
        _check(lib().gi_build_maps(self.handle, ctypes.byref(p)))

    def gather(self, which, queries, k, r, filter=GI_FILTER_DISK, params=None, want_idx=True):
        queries = np.ascontiguousarray(queries, dtype=QUERY)
        n = len(queries)
        idx = np.full((n, k), -1, dtype=np.int32) if want_idx else None
        cnt = np.zeros(n, dtype=np.int32)
        rgb = np.zeros((n, 3), dtype=np.float64)
        p = params or default_params()
        _check(lib().gi_gather_batch(self.handle, which, queries.ctypes.data, n, k, float(r), filter, ctypes.byref(p),
                                     idx.ctypes.data if want_idx else None, cnt.ctypes.data, rgb.ctypes.data))
        return idx, cnt, rgb

    def gather_device(self, which, d_queries, n, k, r, filter, params, d_idx, d_count, d_rgb, stream=None):
        _check(lib().gi_gather_batch_device(self.handle, which, d_queries, n, k, float(r), filter, ctypes.byref(params),
                                            d_idx, d_count, d_rgb, stream))

    def gather_device_noalloc(self, which, d_queries, n, k, r, filter, params, d_idx, d_count, d_rgb, stream=None):
        """gi_gather_batch_device with all scratch in the library's static device array: allocates nothing, n <= NOALLOC_MAX_QUERIES"""
        _check(lib().gi_gather_batch_device_noalloc(self.handle, which, d_queries, n, k, float(r), filter, ctypes.byref(params),
                                                    d_idx, d_count, d_rgb, stream))

    def cached_radiance(self, queries, r):
        """EstimateCachedRadiance per query on the global map (built with params.irradiance_cache): rgb[n, 3]."""
        queries = np.ascontiguousarray(queries, dtype=QUERY)
        rgb = np.zeros((len(queries), 3), dtype=np.float64)
        _check(lib().gi_cached_radiance_batch(self.handle, queries.ctypes.data, len(queries), float(r), rgb.ctypes.data))
        return rgb

    def cached_radiance_device(self, d_queries, n, r, d_rgb, stream=None):
        _check(lib().gi_cached_radiance_batch_device(self.handle, d_queries, n, float(r), d_rgb, stream))

    def trace_photons(self, params):
        """MapPhotons(): returns GiPhotonStats; the maps stay on the device, built."""
        st = GiPhotonStats()
        _check(lib().gi_trace_photons(self.handle, ctypes.byref(params), ctypes.byref(st)))
        return st

    def scale_photon_power(self, which, factor):
        _check(lib().gi_scale_photon_power(self.handle, which, float(factor)))

    def render_float(self, params, width, height, aa=0):
        """The supersampled FP64 frame before clamping: array [W*2^aa, H*2^aa, 3], indexed [i][j] like the
        reference's image_buffer (render.cpp:185)."""
        f = 1 << aa
        out = np.zeros((width * f, height * f, 3), dtype=np.float64)
        st = GiRenderStats()
        _check(lib().gi_render_float(self.handle, ctypes.byref(params), width, height, aa, out.ctypes.data, ctypes.byref(st)))
        return out, st

    def render(self, params, width, height, aa=0, want_image=True):
        """RenderImage(): uint8 array [H, W, 3], row 0 = bottom row (R2Image order).  With a communicator attached the
        frame arrives on rank 0 only; the other ranks pass want_image=False and get None."""
        out = np.zeros((height, width, 3), dtype=np.uint8) if want_image else None
        st = GiRenderStats()
        _check(lib().gi_render(self.handle, ctypes.byref(params), width, height, aa, out.ctypes.data if want_image else None,
                               ctypes.byref(st)))
        return out, st


def parse_flags(argv, params=None):
    """The reference's flag grammar (ParseArgs, src/utils/io_utils.cpp:16-214) applied to a GiParams.
    Returns (params, extra) with extra = {"aa", "width", "height", "real"} for the flags that are not render globals."""
    p = params or default_params()
    extra = {"aa": 2, "width": 1024, "height": 1024, "real": False}
    it = iter(argv)

    def nxt():
        return next(it)
    for a in it:
        if a == "-v": p.verbose = 1
        elif a == "-threads": p.threads = max(1, int(float(nxt())))
        elif a == "-aa": extra["aa"] = abs(int(nxt()))
        elif a == "-real": extra["real"] = True
        elif a == "-no_fresnel": p.fresnel = 0
        elif a == "-ir":
            p.ir_air = float(nxt())
            if p.ir_air <= 0: p.ir_air = 1e-6
        elif a == "-no_ambient": p.ambient = 0
        elif a == "-no_direct": p.direct_illum = 0
        elif a == "-no_transmissive": p.transmissive_illum = 0
        elif a == "-no_specular": p.specular_illum = 0
        elif a == "-no_indirect": p.indirect_illum = 0
        elif a == "-no_caustic": p.caustic_illum = 0
        elif a == "-photon_viz": p.direct_photon_illum = 1
        elif a == "-fast_global": p.fast_global = 1; p.direct_photon_illum = 1; p.indirect_illum = 0
        elif a == "-cache": p.irradiance_cache = 1
        elif a == "-no_monte": p.monte_carlo = 0
        elif a == "-md": p.max_monte_depth = max(1, int(nxt()))
        elif a == "-absorb": p.prob_absorb = max(0.0, float(nxt()))
        elif a == "-no_rs": p.recursive_shadows = 0
        elif a == "-no_dt": p.distrib_transmissive = 0
        elif a == "-tt": p.transmissive_test = max(1, int(nxt()))
        elif a == "-no_ds": p.distrib_specular = 0
        elif a == "-st": p.specular_test = max(1, int(nxt()))
        elif a == "-global": p.global_photon_count = max(1, int(nxt()))
        elif a == "-caustic": p.caustic_photon_count = max(1, int(nxt()))
        elif a == "-pd": p.max_photon_depth = max(1, int(nxt()))
        elif a == "-it": p.indirect_test = max(1, int(nxt()))
        elif a == "-gs": p.global_estimate_size = max(1, int(nxt()))
        elif a == "-gd":
            p.global_estimate_dist = float(nxt())
            if p.global_estimate_dist < 0: p.global_estimate_dist = 1e-6
        elif a in ("-gf", "-cf"):
            kind = nxt()
            f = None
            if kind == "cone":
                f = GI_FILTER_CONE
                p.filter_const_k = float(nxt())
                p.filter_const_k = max(1.0, p.filter_const_k)
            elif kind == "gauss":
                f = GI_FILTER_GAUSS
            if f is not None:
                if a == "-gf": p.global_filter = f
                else: p.caustic_filter = f
        elif a == "-cs": p.caustic_estimate_size = max(1, int(nxt()))
        elif a == "-cd":
            p.caustic_estimate_dist = float(nxt())
            if p.caustic_estimate_dist < 0: p.caustic_estimate_dist = 1e-6
        elif a == "-no_shadow": p.shadows = 0
        elif a == "-no_ss": p.soft_shadows = 0
        elif a == "-lt": p.light_test = max(1, int(nxt()))
        elif a == "-ss": p.shadow_test = max(0, int(nxt()))
        elif a == "-dof":
            p.depth_of_field = 1
            p.dof_test = max(1, int(nxt()))
            p.focus_depth = max(1e-6, float(nxt()))
            p.aperture_radius = float(nxt())
            if p.aperture_radius <= 0: p.aperture_radius = 1e-6
        elif a == "-resolution":
            extra["width"] = abs(int(nxt())); extra["height"] = abs(int(nxt()))
        elif a == "-seed": p.seed = int(nxt())
        else:
            raise ValueError("Invalid program argument: %s" % a)
    return p, extra
