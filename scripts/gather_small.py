"""k = 50 gathers on a 2176-photon map (the global-map regime of a default frame): 2^20 queries at the primary hits of
jensen 1024^2, 4 launches.  Used under ncu (scripts/r02_bench.sh) for profiles/gather_small_map.json."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import torch
import gi_b200
import bench
dev = torch.device("cuda", 0)
ctx = gi_b200.Context(0)
scene = gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", "jensen.gisc"))
ctx.upload_scene(scene)
st = torch.cuda.current_stream().cuda_stream
side = int(sys.argv[1]) if len(sys.argv) > 1 else 1024          # 1024 -> 2^20 queries; 2900 -> the bench section's 2^23
rays = bench.camera_rays(torch, scene.view.camera, side, side, dev)
hits = torch.empty(rays.shape[0] * 72, dtype=torch.uint8, device=dev)
ctx.intersect_device(rays.data_ptr(), rays.shape[0], hits.data_ptr(), st); torch.cuda.synchronize()
q, _, _ = bench.make_queries(torch, rays, hits, bench.scene_tables(torch, scene, dev))
ph = bench.make_photons(torch, ctx, scene, 2176, 7, dev)
ctx.load_photons_device(gi_b200.GI_MAP_GLOBAL, ph.data_ptr(), 2176)
p = gi_b200.default_params(); ctx.build_maps(p)
nq = min(q.shape[0], 1 << 23)
rgb = torch.empty((nq, 3), dtype=torch.float64, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(4):
    if it == 1: e0.record()
    ctx.gather_device(gi_b200.GI_MAP_GLOBAL, q.data_ptr(), nq, int(p.global_estimate_size), float(p.global_estimate_dist), int(p.global_filter), p, None, None, rgb.data_ptr(), st)
e1.record(); torch.cuda.synchronize()
print("small-map gather: queries %d  k %d  %.3f ms per launch  %.1f M gathers/s" % (nq, p.global_estimate_size, e0.elapsed_time(e1) / 3, nq / (e0.elapsed_time(e1) / 3) / 1e3))
