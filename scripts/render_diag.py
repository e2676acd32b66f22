"""Diagnostic: render every fixture case of tests/golden/render.npz on the GPU, print error statistics and save the
frames to gpurun_out/render_diag.npz."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
import test_gpu_render as T

ctx = gi_b200.Context(0)
out = {}
only = sys.argv[1:]
for name in T.DET + T.STO:
    if only and name not in only: continue
    params, w, h = T.setup(ctx, name)
    try:
        img, st = ctx.render_float(params, w, h, 0)
    except Exception as e:
        print(name, "ERROR", e); continue
    ref = T.G[name]
    err = np.abs(img - ref) / np.maximum(1.0, np.abs(ref))
    bad = (err.max(-1) > 1e-9)
    print(f"{name:28s} maxerr {err.max():.3e} bad {bad.sum():5d}/{bad.size} mean ours {img.mean():.5f} ref {ref.mean():.5f} "
          f"rays {st.screen_rays} shadow {st.shadow_rays} monte {st.monte_rays} T {st.transmissive_samples} S {st.specular_samples} "
          f"I {st.indirect_samples} C {st.caustic_samples} ms {st.render_ms:.2f}")
    out[name] = img
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "render_diag.npz"), **out)
