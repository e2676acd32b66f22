#!/usr/bin/env python
"""Generate tests/golden/photons.npz from the REFERENCE's MapPhotons (src/photonmap.cpp:260-436), run through
`oracle/_ref/ref_harness photons` in the build container.

The reference RNG is unseeded, so a traced map cannot be compared photon by photon.  The fixture keeps, per scene and
map, statistics that the photon DISTRIBUTION determines: the number stored, the summed power (sum of decoded RGBE,
equal to total light power x stores per emitted photon), and power / count histograms over an 8x8x8 grid of the scene
box and over the two direction-code bytes.  tests/test_gpu_photons.py compares the GPU tracer's maps to them within
counting noise."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from gi_testlib import *  # noqa

TMP = "/tmp/gi_golden_photons"
CASES = {"cornell": (150000, 150000), "jensen": (150000, 60000), "glass_cube": (100000, 100000), "stack": (150000, 0),
         "softshadow": (100000, 0), "specular": (100000, 50000), "cornell@fast_global": (100000, 0)}


def decode_rgbe(rgbe):
    e = rgbe[:, 3].astype(np.int32)
    inv = np.where(e == 0, 0.0, np.ldexp(1.0, e - 128 - 8))
    return rgbe[:, :3].astype(np.float64) * inv[:, None]


def stats(ph, lo, hi):
    pos = ph["position"]; power = decode_rgbe(ph["rgbe"])
    cell = np.clip(((pos - lo) / np.maximum(hi - lo, 1e-12) * 8).astype(int), 0, 7)
    flat = cell[:, 0] * 64 + cell[:, 1] * 8 + cell[:, 2]
    out = {"n": np.array([len(ph)]), "power": power.sum(0),
           "grid_count": np.bincount(flat, minlength=512).astype(np.float64),
           "grid_power": np.stack([np.bincount(flat, weights=power[:, c], minlength=512) for c in range(3)], 1),
           "phi_hist": np.bincount(ph["direction"] >> 8, minlength=256).astype(np.float64),
           "theta_hist": np.bincount(ph["direction"] & 255, minlength=256).astype(np.float64)}
    return out


def main():
    os.makedirs(TMP, exist_ok=True)
    out = {}
    for case, (ng, nc) in CASES.items():
        scene = case.split("@")[0]
        scn = f"{REF_INPUT}/{scene}.scn"
        args = ["-global", ng, "-threads", 1]
        if case.endswith("@fast_global"):
            args += ["-fast_global", 1]
        args += ["-caustic", nc] if nc else ["-caustic", 1]
        print(case, ref_harness("photons", scn, f"{TMP}/{scene}", *args).stdout.strip())
        tail = np.frombuffer(open(f"{HERE}/{scene}.gisc", "rb").read()[-128:], dtype="<f8")
        lo, hi = tail[6:9], tail[9:12]
        out[f"{case}__target"] = np.array([ng, nc])
        for m in ("global", "caustic"):
            ph = read_photons(f"{TMP}/{scene}.{m}.gipm")
            if m == "caustic" and nc == 0:
                continue
            for k, v in stats(ph, lo, hi).items():
                out[f"{case}__{m}__{k}"] = v
            print(scene, m, len(ph), "power", decode_rgbe(ph["rgbe"]).sum(0))
    np.savez_compressed(f"{HERE}/photons.npz", **out)


if __name__ == "__main__":
    main()
