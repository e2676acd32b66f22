#!/usr/bin/env python
"""Generate tests/golden/parse_args.json: the reference's ParseArgs (src/utils/io_utils.cpp:16-214) run on a list of
argument vectors through `oracle/_ref/ref_harness parse` (build container only)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import subprocess
import numpy as np
from gi_testlib import ref_harness, read_png_rgb, ROOT, REF_INPUT  # noqa

# whole-program runs of the UNMODIFIED reference (oracle/_ref/photonmap_ref) under flag sets that draw no random
# number: name -> (scene, flags).  The decoded PNG pixels go to tests/golden/cli_images.npz.
PROGRAM = {
    "art3": ("art3", "-resolution 64 48 -aa 1 -no_monte -no_transmissive -no_specular -no_indirect -no_caustic"),
    "stack": ("stack", "-resolution 48 48 -aa 2 -no_monte -no_transmissive -no_specular -no_indirect -no_caustic -no_shadow"),
    "specular_md1": ("specular", "-resolution 60 40 -aa 0 -md 1 -no_ds -no_dt -st 1 -tt 1 -no_indirect -no_caustic"),
}

CASES = [
    "",
    "-v -threads 8 -aa 3 -real -no_fresnel -ir 1.3",
    "-threads 0 -aa -1 -ir -2",
    "-no_ambient -no_direct -no_transmissive -no_specular -no_indirect -no_caustic",
    "-photon_viz", "-fast_global", "-no_indirect -fast_global", "-fast_global -no_caustic -cache",
    "-no_monte -md 0 -absorb -1 -no_rs -no_dt -tt 0 -no_ds -st 0",
    "-md 7 -absorb 0.25 -tt 9 -st 11",
    "-global 0 -caustic 0 -pd 0 -it 0 -gs 0 -gd -1 -cs 0 -cd -1",
    "-global 5000 -caustic 123456 -pd 12 -it 64 -gs 100 -gd 0.75 -cs 60 -cd 0.1",
    "-gf cone 0.5", "-gf cone 2.5 -cf gauss", "-cf cone 1.5 -gf gauss", "-gf disk -cf box",
    "-no_shadow -no_ss -lt 0 -ss -5", "-lt 16 -ss 0",
    "-dof 0 0 0", "-dof 16 9.5 0.3", "-resolution -320 240", "-resolution 64 48 -aa 0",
]


def main():
    out = []
    for c in CASES:
        r = ref_harness("parse", *c.split())
        out.append({"argv": c.split(), "globals": json.loads(r.stdout)})
    with open(os.path.join(HERE, "parse_args.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(len(out), "cases")
    imgs = {}
    os.makedirs("/tmp/gi_golden_cli", exist_ok=True)
    for name, (scene, flags) in PROGRAM.items():
        png = f"/tmp/gi_golden_cli/{name}.png"
        subprocess.run([os.path.join(ROOT, "oracle", "_ref", "photonmap_ref"), f"{REF_INPUT}/{scene}.scn", png, *flags.split()],
                       check=True, capture_output=True)
        imgs[name] = read_png_rgb(png)
        imgs[name + "__meta"] = np.array([scene, flags])
        print(name, imgs[name].shape, imgs[name].mean())
    np.savez_compressed(os.path.join(HERE, "cli_images.npz"), **imgs)


if __name__ == "__main__":
    main()
