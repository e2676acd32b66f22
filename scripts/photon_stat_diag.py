"""Diagnostic: the photon fixtures with several seeds under one kernel choice (GI_PHOTON_MEGAKERNEL): bins off per histogram
(the test allows < 2 %), and a checksum of each map (equal checksums under both kernels = identical maps)."""
import os, sys, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
from gi_testlib import GOLDEN
sys.path.insert(0, GOLDEN)
from make_photon_golden import stats
import test_gpu_photons as T
G = T.G
ctx = gi_b200.Context(0)
for case in sys.argv[1:] or ["stack"]:
    for seed in (7, 8, 9, 10):
        params, st, lo, hi = T.trace(ctx, case, seed)
        out = []
        for which, m in ((0, "global"), (1, "caustic")):
            if f"{case}__{m}__n" not in G.files: continue
            ph = ctx.get_photons(which); n_ref = int(G[f"{case}__{m}__n"][0])
            s = stats(ph, lo, hi); ref_power = G[f"{case}__{m}__power"]
            rc = G[f"{case}__{m}__grid_count"]
            def bad(ours, ref):
                fo, fr = ours / max(len(ph), 1), ref / max(n_ref, 1)
                big = ref >= 400
                sigma = np.sqrt(fr / n_ref + fr / len(ph))
                b = np.abs(fo - fr)[big] > 5 * sigma[big] + 0.03 * fr[big]
                return int(b.sum()), int(big.sum())
            res = [bad(s["grid_count"], rc)]
            for c in range(3):
                scale = rc.sum() / max(G[f"{case}__{m}__grid_power"][:, c].sum(), 1e-30)
                res.append(bad(s["grid_power"][:, c] * scale * (ref_power[c] / max(s["power"][c], 1e-30)), G[f"{case}__{m}__grid_power"][:, c] * scale))
            out.append((m, len(ph), zlib.crc32(ph.tobytes()), res))
        print(case, "seed", seed, out)

# which cells, by how much (grid count only)
if os.environ.get("DIAG_CELLS"):
    for seed in (7, 8, 9, 10):
        params, st, lo, hi = T.trace(ctx, "stack", seed)
        ph = ctx.get_photons(0); n_ref = int(G["stack__global__n"][0])
        s = stats(ph, lo, hi); rc = G["stack__global__grid_count"]
        fo, fr = s["grid_count"] / len(ph), rc / n_ref
        big = rc >= 400
        sigma = np.sqrt(fr / n_ref + fr / len(ph))
        dev = (fo - fr) / np.maximum(sigma, 1e-30)
        idx = np.nonzero(big & (np.abs(fo - fr) > 4 * sigma + 0.02 * fr))[0]
        print("seed", seed, [(int(i), int(rc[i]), int(s["grid_count"][i]), round(float(dev[i]), 1)) for i in idx])
