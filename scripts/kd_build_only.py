"""load 1e7 synthetic photons (on the jensen surfaces) and build the map twice; the second build is the one to look at under ncu"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import torch
import gi_b200
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
dev = torch.device("cuda", 0)
ctx = gi_b200.Context(0)
scene = gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", "jensen.gisc"))
ctx.upload_scene(scene)
ph = bench.make_photons(torch, ctx, scene, n, 100, dev)
p = gi_b200.default_params()
for it in range(2):
    ctx.load_photons_device(gi_b200.GI_MAP_CAUSTIC, ph.data_ptr(), n)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ctx.build_maps(p)
    torch.cuda.synchronize(); print("build %d: %.2f ms" % (it, (time.perf_counter() - t0) * 1e3))
