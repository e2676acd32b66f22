"""Diagnostic: stochastic fixture cases, mean of N seeds vs the reference mean."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
import test_gpu_render as T
ctx = gi_b200.Context(0)
runs = int(os.environ.get("RUNS", "32"))
out = {}
for name in T.STO:
    params, w, h = T.setup(ctx, name)
    imgs = []
    for s in range(runs):
        params.seed = 1000 + s
        imgs.append(ctx.render_float(params, w, h, 0)[0])
    imgs = np.stack(imgs); mean = imgs.mean(0); sem = imgs.std(0, ddof=1) / np.sqrt(runs)
    ref, ref_sem, ref_std = T.G[name], T.G[name + "__sem"], T.G[name + "__std"]
    z = np.abs(mean - ref) / (np.sqrt(sem ** 2 + ref_sem ** 2) + 1e-4)
    print(f"{name:22s} mean ours {mean.mean(axis=(0,1))} ref {ref.mean(axis=(0,1))} pix z>4: {np.mean(z > 4):.3f} zmax {z.max():.1f} "
          f"std ours {imgs.std(0, ddof=1).mean():.4f} ref {ref_std.mean():.4f}")
    out[name] = mean; out[name + "__sem"] = sem
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "render_stat.npz"), **out)
