import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
import test_gpu_render as T
ctx = gi_b200.Context(0)
T.setup(ctx, "jensen_noshadow")
for lt in (1, 4, 16, 64):
    params, _ = gi_b200.parse_flags(f"-no_monte -no_shadow -lt {lt}".split())
    imgs = []
    for s in range(24):
        params.seed = 1000 + s
        imgs.append(ctx.render_float(params, 24, 24, 0)[0])
    imgs = np.stack(imgs)
    print(lt, imgs.std(0, ddof=1).mean(), imgs.mean(), "seed-to-seed identical:", np.array_equal(imgs[0], imgs[1]))
