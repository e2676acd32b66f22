#!/usr/bin/env python
"""Generate tests/golden/render.npz from the REFERENCE ITSELF (run in the build container only).

For every case below the reference's own RayTrace (src/raytracer.cpp:174-233) is run over the reference's own camera
rays (src/render.cpp:64-117) through `oracle/_ref/ref_harness shade`, which parses the flags with the reference's
ParseArgs.  The result is image_buffer[i][j] before clamping, the thing gi_render_float returns.

  deterministic cases  flag sets under which RayTrace draws no random number that can change a pixel (point / spot /
                       directional lights, hard shadows, Monte Carlo paths of depth 1 without importance sampling,
                       photon-map lookups at the primary hit).  tests compare to 1e-9.
  stochastic cases     soft shadows, distributed reflection / refraction, indirect sampling.  The reference RNG is
                       unseeded (random_device), so the fixture holds the MEAN of several reference runs plus the
                       per-pixel standard error; tests compare block averages within that error.
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa

TMP = "/tmp/gi_golden_render"

# name -> (scene, W, H, flags, photon maps to load)
DET = {
    "cornell_direct": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular", ()),
    "art3_direct": ("art3", 40, 30, "-no_monte -no_transmissive -no_specular", ()),
    "stack_direct": ("stack", 40, 30, "-no_monte -no_transmissive -no_specular", ()),
    "transform_direct": ("transform", 40, 30, "-no_monte -no_transmissive -no_specular", ()),
    "specular_noshadow": ("specular", 40, 30, "-no_monte -no_shadow", ()),
    "glass_cube_direct": ("glass_cube", 32, 32, "-no_monte", ()),
    "stilllife_noambient": ("stilllife", 32, 24, "-no_monte -no_ambient", ()),
    # depth-1 paths, exact bounce directions: reflective + transmissive bounce, Fresnel split, in-path direct light
    "specular_md1": ("specular", 40, 30, "-md 1 -no_ds -no_dt -st 1 -tt 1", ()),
    "glass_cube_md1": ("glass_cube", 32, 32, "-md 1 -no_ds -no_dt -st 1 -tt 1", ()),
    "glass_cube_md1_nofresnel": ("glass_cube", 32, 32, "-md 1 -no_ds -no_dt -st 1 -tt 1 -no_fresnel -ir 1.2", ()),
    "stilllife_md1": ("stilllife", 32, 24, "-md 1 -no_ds -no_dt -st 2 -tt 3", ()),
    "cornell_md1": ("cornell", 40, 40, "-md 1 -no_ds -no_dt -st 1 -tt 1 -no_rs", ()),
    # photon-map lookups at the primary hit
    "cornell_caustic": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular -no_indirect -no_direct", ("caustic",)),
    "cornell_viz": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular -no_indirect -no_caustic -photon_viz -gs 30 -gd 1.5",
                    ("global",)),
    "cornell_viz_cone": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular -no_indirect -photon_viz -gf cone 1.2 -cf gauss",
                         ("global", "caustic")),
    # irradiance cache (-cache): pre-pass on the loaded global map, EstimateCachedRadiance at the primary hit
    "cornell_viz_cache": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular -no_indirect -no_caustic -photon_viz -cache",
                          ("global",)),
    "cornell_viz_cache_gs": ("cornell", 40, 40, "-no_monte -no_transmissive -no_specular -no_indirect -photon_viz -cache -gs 20 -gd 1.0",
                             ("global", "caustic")),
    # depth-1 paths with caustic lookups at the path vertex and the fast-global estimate on a diffuse pick is random:
    # keep it out of the deterministic set
}

# name -> (scene, W, H, flags, maps, reference runs)
STO = {
    "softshadow_soft": ("softshadow", 24, 24, "-no_monte -lt 16 -ss 16", (), 64),
    "jensen_soft": ("jensen", 24, 24, "-no_monte -no_transmissive -no_specular -lt 8 -ss 8", (), 64),
    "jensen_noshadow": ("jensen", 24, 24, "-no_monte -no_shadow -lt 16", (), 64),
    "specular_distrib": ("specular", 24, 18, "-md 4 -st 8 -tt 8", (), 64),
    "glass_cube_paths": ("glass_cube", 24, 24, "-md 8 -st 8 -tt 8", (), 64),
    "cornell_indirect": ("cornell", 24, 24, "-no_transmissive -no_specular -no_caustic -it 16 -md 6", ("global",), 64),
    "cornell_full": ("cornell", 24, 24, "-it 8 -st 4 -tt 4 -md 6", ("global", "caustic"), 64),
    "cornell_fast_global": ("cornell", 24, 24, "-fast_global -st 8 -tt 8 -md 6", ("global", "caustic"), 64),
    "cornell_indirect_cache": ("cornell", 24, 24, "-no_transmissive -no_specular -no_caustic -it 16 -md 6 -cache", ("global",), 64),
    "cornell_fast_global_cache": ("cornell", 24, 24, "-fast_global -cache -st 8 -tt 8 -md 6", ("global", "caustic"), 64),
    # EVERY flag at its default (lt / ss / tt / st 128, it 256, md 128, k 50 / 225): the sample counts of the metric's frame.
    # 32 reference runs; the fixture also keeps the two half-means so that the test can state a PSNR against the
    # reference's own run-to-run noise floor.
    "cornell_defaults": ("cornell", 48, 48, "", ("global", "caustic"), 32),
}


def shade(scene, w, h, flags, maps, tag, dof=False):
    scn = f"{REF_INPUT}/{scene}.scn"
    rays = f"{TMP}/{scene}.{w}x{h}.rays.bin"
    if not os.path.exists(rays):
        ref_harness("camrays", scn, w, h, rays)
    args = []
    for m in maps:
        args += [f"-load_{m}", f"{HERE}/cornell.{m}.gipm"]
    out = f"{TMP}/{tag}.bin"
    ref_harness("shade", scn, rays, out, *args, *flags.split(), "-threads", 8)
    raw = open(out, "rb").read()
    n = int(np.frombuffer(raw[4:8], dtype="<u4")[0])
    return np.frombuffer(raw[8:], dtype="<f8").reshape(n, 3).reshape(w, h, 3).copy()


def main():
    os.makedirs(TMP, exist_ok=True)
    out = {}
    only = None
    if len(sys.argv) > 2 and sys.argv[1] == "--only":          # add / refresh the named cases, keep the rest of the file
        only = set(sys.argv[2].split(","))
        out = dict(np.load(f"{HERE}/render.npz"))
    for name, (scene, w, h, flags, maps) in DET.items():
        if only is not None and name not in only:
            continue
        a = shade(scene, w, h, flags, maps, name)
        b = shade(scene, w, h, flags, maps, name + ".again")
        # (photon sums may differ in the last bits from run to run: the reference kd-tree returns its k nearest in a
        # run-dependent order)
        assert np.allclose(a, b, rtol=1e-12, atol=1e-13), f"{name}: the reference is not deterministic under these flags"
        out[name] = a
        out[name + "__meta"] = np.array([scene, str(w), str(h), flags, ",".join(maps)])
        print(name, "mean", a.mean(axis=(0, 1)))
    for name, (scene, w, h, flags, maps, runs) in STO.items():
        if only is not None and name not in only:
            continue
        imgs = np.stack([shade(scene, w, h, flags, maps, f"{name}.{r}") for r in range(runs)])
        out[name] = imgs.mean(0)
        out[name + "__sem"] = imgs.std(0, ddof=1) / np.sqrt(runs)
        out[name + "__std"] = imgs.std(0, ddof=1)
        if name == "cornell_defaults":
            out[name + "__half_a"] = imgs[: runs // 2].mean(0); out[name + "__half_b"] = imgs[runs // 2:].mean(0)
        out[name + "__meta"] = np.array([scene, str(w), str(h), flags, ",".join(maps)])
        print(name, "mean", imgs.mean(axis=(0, 1, 2)), "max sem", out[name + "__sem"].max())
    np.savez_compressed(f"{HERE}/render.npz", **out)


if __name__ == "__main__":
    main()
