#!/usr/bin/env python
"""bench.py -- the hot path of the photon-mapping renderer on B200: closest-hit ray casting + photon kNN gather.

One "step" = one pass of the hot path over one frame's worth of synthetic input of BASELINE.json configs[1]
(jensen scene, 1024 x 1024, aa 2  =>  4096 x 4096 = 16.8 M primary rays): gi_intersect_batch over every camera ray,
then gi_gather_batch (k = 225, r = 0.225, the reference's caustic defaults) at every primary hit point against a
10^7-photon map (the reference's default caustic map size) laid on the scene's surfaces.
metric = Mrays/s in the reference's own accounting (render.cpp:224-255 counts gather samples as rays), with the
two stages also reported separately (mrays_per_s_intersect, gathers_per_s).

  python bench.py [--gpus N --steps K --warmup W]          our arm (N > 1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      the reference's own CPU code (oracle/_ref) on the host cores

`value` is measured with all inputs resident in HBM; `e2e` goes through the host-buffer C-ABI calls (pinned host
memory in, host results out, copies inside the timed region).  torch is used for device buffers, events and
torch.distributed only; every kernel in the timed region is ours (libgi_b200.so).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLDEN = os.path.join(ROOT, "tests", "golden")

SCENE = "jensen"
K, R, FILTER = 225, 0.225, 0
B_PHOTON = 32          # bytes per photon touched in the parity layout: 24 B FP64 position + 8 B shade record
GATHER_ALGO_BYTES = K * B_PHOTON + 52      # k x B_photon + query in / rgb out
RAY_ALGO_BYTES = 48 + 72                   # GiRay in + GiHit out


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--aa", type=int, default=2)
    ap.add_argument("--photons", type=int, default=10_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-rays", type=int, default=2_000_000)
    ap.add_argument("--cpu-queries", type=int, default=200_000)
    ap.add_argument("--no-frame", action="store_true", help="skip the whole-program frame (photon pass + render) report")
    ap.add_argument("--frame-scene", default="cornell")
    ap.add_argument("--frame-res", type=int, default=512)
    ap.add_argument("--frame-aa", type=int, default=1)
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index; self.proc = None; self.path = None

    def start(self):
        self.path = tempfile.mktemp(suffix=".csv")
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons, mx = [], set(), None
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.unlink(self.path)
        except OSError:
            pass
        return out


def camera_rays(torch, cam, W, H, dev, jitter=0.0):
    """render.cpp:64-117 without depth of field: far_point = eye + towards*F + right*tan(xfov)*F*dx + up*tan(yfov)*F*dy,
    dx = 2(i - W/2)/W, dy = 2(j - H/2)/H, rays normalised (R3Ray ctor).  Column-major like image_buffer[i][j]."""
    f64 = torch.float64
    eye = torch.tensor(list(cam.eye), dtype=f64, device=dev)
    tow = torch.tensor(list(cam.towards), dtype=f64, device=dev)
    up = torch.tensor(list(cam.up), dtype=f64, device=dev)
    right = torch.tensor(list(cam.right), dtype=f64, device=dev)
    F = 100.0
    i = torch.arange(W, dtype=f64, device=dev) + jitter
    j = torch.arange(H, dtype=f64, device=dev) + jitter
    dx = (2.0 * (i - (W // 2)) / W)[:, None, None]
    dy = (2.0 * (j - (H // 2)) / H)[None, :, None]
    far = eye + tow * F + right * (np.tan(cam.xfov) * F) * dx + up * (np.tan(cam.yfov) * F) * dy
    d = far - eye
    d = d / torch.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]))[..., None]
    rays = torch.empty((W, H, 6), dtype=f64, device=dev)
    rays[..., 0:3] = eye
    rays[..., 3:6] = d
    return rays.reshape(-1, 6).contiguous()


def hits_fields(torch, hits_u8):
    """view a GiHit[n] byte tensor as (hit int32[n], doubles [n, 9] from byte 0; columns 2..8 = t, p, n)"""
    n = hits_u8.shape[0] // 72
    as_i32 = hits_u8.view(torch.int32).reshape(n, 18)
    as_f64 = hits_u8.view(torch.float64).reshape(n, 9)
    return as_i32[:, 0], as_f64


def make_queries(torch, rays, hits_u8):
    """GiGatherQuery records at the hit points, set up as CausticIllumination is called from RayTrace
    (raytracer.cpp:138-149, 174-190): cos_theta = n.(-view), exact_bounce = mirror direction."""
    hit, f = hits_fields(torch, hits_u8)
    m = hit == 1
    d = rays[m, 3:6]; p = f[m, 3:6]; n = f[m, 6:9]
    cos = -(n * d).sum(1)
    q = torch.zeros((int(m.sum()), 17), dtype=torch.float64, device=rays.device)
    q[:, 0:3] = p; q[:, 3:6] = n; q[:, 6:9] = d + 2.0 * cos[:, None] * n; q[:, 9] = cos
    q[:, 10:13] = torch.tensor([0.5, 0.5, 0.5], dtype=torch.float64, device=rays.device)
    q[:, 13:16] = 0.0; q[:, 16] = 1.0
    return q.contiguous()


def make_photons(torch, ctx, scene, n, seed, dev):
    """synthetic photon map: n photons at the closest hits of uniformly random rays, i.e. spread over the scene's
    surfaces like real photons; random power and incident-direction codes"""
    g = torch.Generator(device=dev); g.manual_seed(seed)
    lo = torch.tensor(list(scene.view.bbox_min), dtype=torch.float64, device=dev)
    hi = torch.tensor(list(scene.view.bbox_max), dtype=torch.float64, device=dev)
    out = torch.zeros((n, 32), dtype=torch.uint8, device=dev)
    filled = 0
    while filled < n:
        m = min(int((n - filled) * 1.3) + 1024, 8_000_000)
        o = lo + torch.rand((m, 3), generator=g, dtype=torch.float64, device=dev) * (hi - lo)
        d = torch.randn((m, 3), generator=g, dtype=torch.float64, device=dev)
        d = d / torch.sqrt((d * d).sum(1))[:, None]
        rays = torch.cat([o, d], dim=1).contiguous()
        hits = torch.empty(m * 72, dtype=torch.uint8, device=dev)
        ctx.intersect_device(rays.data_ptr(), m, hits.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        hit, f = hits_fields(torch, hits)
        p = f[hit == 1, 3:6][: n - filled]
        c = p.shape[0]
        rec = out[filled:filled + c]
        rec[:, 0:24] = p.contiguous().view(torch.uint8).reshape(c, 24)
        rec[:, 24:27] = torch.randint(1, 256, (c, 3), generator=g, dtype=torch.uint8, device=dev)
        rec[:, 27] = 120
        rec[:, 28:30] = torch.randint(0, 256, (c, 2), generator=g, dtype=torch.uint8, device=dev)
        filled += c
    return out


# ------------------------------------------------------------------------------------------- reference / CPU arm
REF_HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
SCENE_TOOL = os.path.join(ROOT, "global-illumination_b200", "bin", "gi_scene_tool")


def export_scene_scn(tmp):
    """the reference binary reads .scn only: write the flattened jensen scene back out as Princeton text"""
    scn = os.path.join(tmp, SCENE + ".scn")
    subprocess.run([SCENE_TOOL, "export", os.path.join(GOLDEN, SCENE + ".gisc"), scn], check=True)
    return scn


def ref_bench(tmp, scn, rays, photons_u8, queries, steps, warmup, threads):
    """time the reference's R3Scene::Intersects + EstimateRadiance (oracle/_ref/ref_harness bench) on host threads"""
    from gi_testlib import write_rays, write_queries, GI_QUERY
    write_rays(os.path.join(tmp, "rays.bin"), rays)
    q = np.zeros(len(queries), dtype=GI_QUERY)
    q.view(np.float64).reshape(-1, 17)[:] = queries
    write_queries(os.path.join(tmp, "q.bin"), q)
    with open(os.path.join(tmp, "ph.gipm"), "wb") as f:
        f.write(np.array([0x4D504947, 1], dtype="<u4").tobytes())
        f.write(np.array([len(photons_u8) // 32], dtype="<u8").tobytes())
        f.write(photons_u8.tobytes())
    out = subprocess.run([REF_HARNESS, "bench", scn, os.path.join(tmp, "rays.bin"), os.path.join(tmp, "ph.gipm"),
                          os.path.join(tmp, "q.bin"), str(K), str(R), str(FILTER), str(threads), str(steps), str(warmup)],
                         check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def cpu_inputs_numpy(args, tmp, scn, n_rays, n_queries, threads):
    """inputs of the same workload generated WITHOUT the GPU (reference arm): camera-ray sample, photons at the
    reference's own closest hits of random rays, queries at the reference's primary hits"""
    from gi_testlib import write_rays, read_ref_hits, random_rays
    import gi_b200
    scene = gi_b200.HostScene(os.path.join(GOLDEN, SCENE + ".gisc"))
    cam = scene.view.camera
    W, H = args.width << args.aa, args.height << args.aa
    rng = np.random.default_rng(1)
    pick = rng.choice(W * H, size=min(n_rays, W * H), replace=False)
    i = (pick // H).astype(np.float64); j = (pick % H).astype(np.float64)
    eye = np.array(cam.eye[:]); tow = np.array(cam.towards[:]); up = np.array(cam.up[:]); right = np.array(cam.right[:])
    F = 100.0
    dx = (2.0 * (i - (W // 2)) / W)[:, None]; dy = (2.0 * (j - (H // 2)) / H)[:, None]
    far = eye + tow * F + right * (np.tan(cam.xfov) * F) * dx + up * (np.tan(cam.yfov) * F) * dy
    d = far - eye
    d /= np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])[:, None]
    rays = np.concatenate([np.broadcast_to(eye, d.shape), d], axis=1)

    def ref_isect(r):
        write_rays(os.path.join(tmp, "gen.bin"), r)
        subprocess.run([REF_HARNESS, "isect", scn, os.path.join(tmp, "gen.bin"), os.path.join(tmp, "gen.hits"), str(threads)],
                       check=True, capture_output=True)
        return read_ref_hits(os.path.join(tmp, "gen.hits"))

    h = ref_isect(rays[:n_queries])
    m = h["hit"] == 1
    dd = rays[:n_queries][m, 3:6]; n = h["n"][m]; cos = -(n * dd).sum(1)
    q = np.zeros((m.sum(), 17))
    q[:, 0:3] = h["p"][m]; q[:, 3:6] = n; q[:, 6:9] = dd + 2.0 * cos[:, None] * n; q[:, 9] = cos
    q[:, 10:13] = 0.5; q[:, 16] = 1.0
    lo = np.array(scene.view.bbox_min[:]); hi = np.array(scene.view.bbox_max[:])
    ph = np.zeros((args.photons, 32), dtype=np.uint8); filled = 0; seed = 50
    while filled < args.photons:
        rr = random_rays(min(int((args.photons - filled) * 1.3) + 1024, 4_000_000), lo, hi, seed, scale=1.0); seed += 1
        hh = ref_isect(rr)
        p = hh["p"][hh["hit"] == 1][: args.photons - filled]
        c = len(p)
        ph[filled:filled + c, 0:24] = np.ascontiguousarray(p).view(np.uint8).reshape(c, 24)
        ph[filled:filled + c, 24:27] = rng.integers(1, 256, (c, 3)); ph[filled:filled + c, 27] = 120
        ph[filled:filled + c, 28:30] = rng.integers(0, 256, (c, 2))
        filled += c
    return rays, ph.reshape(-1), q


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if not os.path.exists(REF_HARNESS):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_harness was not built (no reference tree at build time)"}))
        return
    threads = os.cpu_count() or 1
    with tempfile.TemporaryDirectory() as tmp:
        scn = export_scene_scn(tmp)
        rays, ph, q = cpu_inputs_numpy(args, tmp, scn, args.cpu_rays, args.cpu_queries, threads)
        r = ref_bench(tmp, scn, rays, ph, q, args.steps, args.warmup, threads)
    sec = r["isect_seconds_per_step"] + r["gather_seconds_per_step"]
    units = len(rays) + len(q)
    val = units / sec / 1e6
    sample = (f"{len(rays)} camera rays + {len(q)} gathers (k={K}, r={R}) per step, a uniform sample of the "
              f"{args.width}x{args.height} aa {args.aa} frame; {args.photons}-photon map; kd-tree build "
              f"({r['kd_build_seconds']:.1f} s, single-threaded) not timed")
    line = {"impl": "reference", "metric": "Mrays/s (closest-hit rays + photon kNN gather samples, reference accounting)",
            "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(args),
            "mrays_per_s_intersect": len(rays) / r["isect_seconds_per_step"] / 1e6,
            "gathers_per_s": len(q) / r["gather_seconds_per_step"],
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def config_dict(args):
    return {"workload": f"input/{SCENE}.scn {args.width}x{args.height} -aa {args.aa}: closest hit for all "
                        f"{(args.width << args.aa) * (args.height << args.aa)} primary rays + caustic-style photon gather "
                        f"(k={K}, r={R}, disk filter) at every primary hit, {args.photons}-photon map on the scene surfaces",
            "photons": args.photons, "k": K, "r": R, "l2": "inputs larger than L2 (0.8 GB rays, 2.3 GB queries per step)"}


# ------------------------------------------------------------------------------------------- our arm (B200)
def run_ours(args):
    import torch
    import torch.distributed as dist
    import gi_b200

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ctx = gi_b200.Context(local_rank)
    scene = gi_b200.HostScene(os.path.join(GOLDEN, SCENE + ".gisc"))
    ctx.upload_scene(scene)
    work_stream = torch.cuda.Stream(device=dev)          # a real (non-default) stream: kernels and events both go here
    torch.cuda.set_stream(work_stream)
    stream = work_stream.cuda_stream
    W, H = args.width << args.aa, args.height << args.aa
    n_rays = W * H
    # weak scaling: every rank renders its own full frame (sub-pixel offset per rank so the inputs differ)
    rays = camera_rays(torch, scene.view.camera, W, H, dev, jitter=0.01 * rank)
    hits = torch.empty(n_rays * 72, dtype=torch.uint8, device=dev)

    # ---- photon map: each rank lays down photons/world photons, one NCCL all-gather replicates the map
    per_rank = args.photons // world
    local_ph = make_photons(torch, ctx, scene, per_rank, 100 + rank, dev)
    allgather_ms = 0.0
    if world > 1:
        full = torch.empty((per_rank * world, 32), dtype=torch.uint8, device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier(); torch.cuda.synchronize()
        e0.record(); dist.all_gather_into_tensor(full, local_ph); e1.record(); torch.cuda.synchronize()
        allgather_ms = e0.elapsed_time(e1)
    else:
        full = local_ph
    n_ph = full.shape[0]
    ctx.load_photons_device(gi_b200.GI_MAP_CAUSTIC, full.data_ptr(), n_ph)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ctx.build_maps()
    torch.cuda.synchronize(); kd_build_ms = (time.perf_counter() - t0) * 1e3

    ctx.intersect_device(rays.data_ptr(), n_rays, hits.data_ptr(), stream)
    torch.cuda.synchronize()
    queries = make_queries(torch, rays, hits)
    n_q = queries.shape[0]
    rgb = torch.empty((n_q, 3), dtype=torch.float64, device=dev)
    cnt = torch.empty(n_q, dtype=torch.int32, device=dev)
    params = gi_b200.default_params()

    def step():
        ctx.intersect_device(rays.data_ptr(), n_rays, hits.data_ptr(), stream)
        ctx.gather_device(gi_b200.GI_MAP_CAUSTIC, queries.data_ptr(), n_q, K, R, FILTER, params, None, cnt.data_ptr(),
                          rgb.data_ptr(), stream)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 1)]
    launches0 = ctx.launch_count()
    torch.cuda.synchronize()
    ev[0].record()
    for s in range(args.steps):
        ctx.intersect_device(rays.data_ptr(), n_rays, hits.data_ptr(), stream)
        ev[2 * s + 1].record()
        ctx.gather_device(gi_b200.GI_MAP_CAUSTIC, queries.data_ptr(), n_q, K, R, FILTER, params, None, cnt.data_ptr(),
                          rgb.data_ptr(), stream)
        ev[2 * s + 2].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launch_count() - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    isect_ms = sum(ev[2 * s].elapsed_time(ev[2 * s + 1]) for s in range(args.steps)) / args.steps
    gather_ms = sum(ev[2 * s + 1].elapsed_time(ev[2 * s + 2]) for s in range(args.steps)) / args.steps
    found_full = int((cnt == K).sum().item())

    # ---- end to end through the host-buffer C ABI: pinned host inputs, H2D + kernels + D2H inside the timed region
    rays_h = torch.empty((n_rays, 6), dtype=torch.float64, pin_memory=True); rays_h.copy_(rays)
    q_h = torch.empty((n_q, 17), dtype=torch.float64, pin_memory=True); q_h.copy_(queries)
    hits_h = torch.empty(n_rays * 72, dtype=torch.uint8, pin_memory=True)
    rgb_h = torch.empty((n_q, 3), dtype=torch.float64, pin_memory=True)
    cnt_h = torch.empty(n_q, dtype=torch.int32, pin_memory=True)
    L = gi_b200.lib()

    def e2e_step():
        gi_b200._check(L.gi_intersect_batch(ctx.handle, rays_h.data_ptr(), n_rays, hits_h.data_ptr()))
        gi_b200._check(L.gi_gather_batch(ctx.handle, gi_b200.GI_MAP_CAUSTIC, q_h.data_ptr(), n_q, K, R, FILTER,
                                         ctypes.byref(params), None, cnt_h.data_ptr(), rgb_h.data_ptr()))
    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps

    # ---- max over ranks
    t = torch.tensor([total_ms, isect_ms, gather_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    units = torch.tensor([float(n_rays), float(n_q)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(units, op=dist.ReduceOp.SUM)
    total_ms, isect_ms, gather_ms, e2e_ms = t.tolist()
    rays_all, q_all = units.tolist()
    ms_per_step = total_ms / args.steps

    frame = None
    if not args.no_frame:
        del rays_h, q_h, hits_h, rgb_h, cnt_h
        frame = frame_report(args, torch, dist, ctx, dev, rank, world)

    if rank == 0:
        peak, peak_src = load_peaks()
        achieved = (n_q * GATHER_ALGO_BYTES) / (gather_ms * 1e-3) / 1e9
        line = {"metric": "Mrays/s (closest-hit rays + photon kNN gather samples, reference accounting)",
                "value": (rays_all + q_all) / (ms_per_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args),
                "mrays_per_s_intersect": rays_all / (isect_ms * 1e-3) / 1e6, "gathers_per_s": q_all / (gather_ms * 1e-3),
                "intersect_ms": isect_ms, "gather_ms": gather_ms, "kd_build_ms": kd_build_ms, "photon_allgather_ms": allgather_ms,
                "photons_in_map": n_ph, "queries_with_k_photons": found_full,
                "roofline": {"kernel": "gather_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                             "algorithmic_bytes_per_gather": GATHER_ALGO_BYTES,
                             "intersect_kernel_GBps": (n_rays * RAY_ALGO_BYTES) / (isect_ms * 1e-3) / 1e9},
                "e2e": {"value": (rays_all + q_all) / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                        "h2d_bytes_per_step": n_rays * 48 + n_q * 136, "d2h_bytes_per_step": n_rays * 72 + n_q * 28,
                        "ms_per_step": e2e_ms},
                "gpu_launches": int(launches), "clocks": clocks, "frame": frame}
        traffic_file = os.path.join(ROOT, "profiles", "gather_traffic.json")
        if os.path.exists(traffic_file):       # ncu capture at a smaller launch: DRAM bytes per gather x this launch's gathers
            tf = json.load(open(traffic_file))
            line["roofline"]["traffic"] = tf["dram_bytes_per_launch"] / tf["queries_per_launch"] * n_q
            line["roofline"]["traffic_note"] = "ncu dram bytes per gather (%s) scaled to this launch" % tf["source"]
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, torch, rays, full, queries)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def frame_report(args, torch, dist, ctx, dev, rank, world):
    """Whole-program frame outside the timed steps: BASELINE.json configs[0] (cornell 512 x 512 -aa 1, every flag at its
    default: 2176 global + 10^7 caustic photons, k 50 / 225, lt/ss/tt/st 128, it 256) -- MapPhotons + RenderImage
    through gi_trace_photons / gi_render, sharded over the ranks when N > 1 (photons per rank + one all-gather, pixel
    columns per rank + one reduce).  Times are wall clock around synchronised calls, max over ranks."""
    import gi_b200
    import multi_gpu
    scene = gi_b200.HostScene(os.path.join(GOLDEN, args.frame_scene + ".gisc"))
    ctx.upload_scene(scene)
    W = H = args.frame_res
    split = None
    for attempt in range(2):        # the first pass pays one-time costs (workspace allocation, NCCL channel set-up); report the second
        params, _ = gi_b200.parse_flags(["-seed", "1"])
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        if world > 1:
            totals = multi_gpu.trace_photons_sharded(ctx, params, dist, torch, dev)
            stored = (totals["global_stored"], totals["caustic_stored"])
        else:
            st = ctx.trace_photons(params)
            params.indirect_illum, params.caustic_illum, params.direct_photon_illum = st.indirect_illum, st.caustic_illum, st.direct_photon_illum
            stored = (int(st.global_stored), int(st.caustic_stored))
            split = {"trace_ms": float(st.trace_ms), "kd_build_ms": float(st.build_ms)}
        torch.cuda.synchronize(); t1 = time.perf_counter()
        if world > 1:
            img, rs = multi_gpu.render_sharded(ctx, params, W, H, args.frame_aa, dist, torch, dev)
        else:
            img, rs = ctx.render(params, W, H, args.frame_aa)
        torch.cuda.synchronize(); t2 = time.perf_counter()
    names = ["screen_rays", "shadow_rays", "monte_rays", "transmissive_samples", "specular_samples", "indirect_samples", "caustic_samples"]
    c = torch.tensor([float(getattr(rs, n)) for n in names], dtype=torch.float64, device=dev)
    t = torch.tensor([(t1 - t0) * 1e3, (t2 - t1) * 1e3, rs.render_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(c, op=dist.ReduceOp.SUM); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    counters = dict(zip(names, [int(x) for x in c.tolist()]))
    photon_ms, render_ms, render_kernel_ms = t.tolist()
    total = sum(counters.values())
    return {"workload": f"input/{args.frame_scene}.scn {W}x{H} -aa {args.frame_aa}, all flags default (BASELINE.json configs[0]): "
                        "gi_trace_photons + gi_render, u8 frame back on the host",
            "photon_pass_ms": photon_ms, "photon_pass_split": split, "render_ms": render_ms, "frame_ms": photon_ms + render_ms, "render_device_ms": render_kernel_ms,
            "photons_stored": {"global": stored[0], "caustic": stored[1]}, "ray_ops": total, "counters": counters,
            "mray_ops_per_s": total / (render_ms * 1e-3) / 1e6,
            "gathers_per_s": (counters["indirect_samples"] + counters["caustic_samples"]) / (render_ms * 1e-3),
            "image_mean": float(img.mean()) if img is not None else None, "timing": "wall clock of the second of two passes, max over ranks; not part of the timed steps"}


def cpu_baseline(args, torch, rays, photons, queries):
    """the reference's own code (oracle/_ref, R3Scene::Intersects + EstimateRadiance) on all host cores, on a bounded
    uniform sample of the same step; falls back to the oracle restatement where the reference was not built"""
    threads = os.cpu_count() or 1
    g = torch.Generator(device="cpu"); g.manual_seed(3)
    ri = torch.randperm(rays.shape[0], generator=g)[: args.cpu_rays].to(rays.device)
    qi = torch.randperm(queries.shape[0], generator=g)[: args.cpu_queries].to(rays.device)
    r = rays[ri].cpu().numpy(); q = queries[qi].cpu().numpy(); ph = photons.reshape(-1).cpu().numpy()
    sample = (f"{len(r)} camera rays + {len(q)} gathers (k={K}, r={R}), uniform sample of the step; same "
              f"{len(ph) // 32}-photon map; kd-tree build not timed")
    with tempfile.TemporaryDirectory() as tmp:
        if os.path.exists(REF_HARNESS):
            res = ref_bench(tmp, export_scene_scn(tmp), r, ph, q, 1, 0, threads)
            sec = res["isect_seconds_per_step"] + res["gather_seconds_per_step"]
            return {"value": (len(r) + len(q)) / sec / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference",
                    "sample": sample, "mrays_per_s_intersect": len(r) / res["isect_seconds_per_step"] / 1e6,
                    "gathers_per_s": len(q) / res["gather_seconds_per_step"], "kd_build_seconds": res["kd_build_seconds"]}
    from gi_testlib import Oracle, GI_PHOTON, GI_QUERY
    o = Oracle()
    h = o.load_scene(os.path.join(GOLDEN, SCENE + ".gisc"))
    q = q[:2000]
    t0 = time.perf_counter(); o.intersect(h, r, threads=threads); t1 = time.perf_counter()
    qq = np.zeros(len(q), dtype=GI_QUERY); qq.view(np.float64).reshape(-1, 17)[:] = q
    o.gather(ph.view(GI_PHOTON), qq, K, R, FILTER, threads=threads, want_idx=False); t2 = time.perf_counter()
    return {"value": (len(r) + len(q)) / (t2 - t0) / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
            "sample": sample + " (brute-force oracle gather, 2000 queries)"}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
