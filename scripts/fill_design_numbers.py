"""Rewrites the measured blocks of DESIGN.md (section 5 table, per-stage sentence, section 6 scaling table + configs paragraph)
and the numbers paragraph of README.md from profiles/r02_bench_n{1,2,4,8}.json and profiles/r02_bench_reference_arm.json."""
import json, os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = lambda *a: os.path.join(ROOT, *a)
B = {n: json.loads(open(P("profiles", f"r02_bench_n{n}.json")).read().strip().splitlines()[-1]) for n in (1, 2, 4, 8)}
R = json.loads(open(P("profiles", "r02_bench_reference_arm.json")).read().strip().splitlines()[-1])
d = B[1]; f1 = d["frames"]["cornell_1024_aa2"]; f5 = d["frames"]["cornell_512_aa1"]
f8 = dict(json.load(open(P("profiles", "r02_frames_all_n8.json")))["frames"]); f8.update(B[8]["frames"])      # configs 3-5 on 8 GPUs: the --frames all run
st = f1["stage_ms_max_over_ranks"]
stages = ", ".join(f"{k[:-3].replace('_', ' ')} {v:.0f}" for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v >= 1.0) + " ms"
table = f"""| | this repo | reference, 16 host cores |
|---|---|---|
| step, 1 × B200: `value` | **{d['value']:.0f} Mrays/s** ({d['ms_per_step']:.1f} ms: replicate {d['stage_ms']['photon_replicate']:.1f}, kd build {d['stage_ms']['kd_build']:.1f}, intersect {d['stage_ms']['intersect']:.2f}, gather {d['stage_ms']['gather']:.1f}, resolve + strips {d['stage_ms']['resolve_glue_and_strip_gather']:.1f}) — round-1 definition (intersect + gather only): {d['value_render_stages_only']:.0f} Mrays/s (372 in round 1) | {R['value']:.3f} Mrays/s (closest hit {R['mrays_per_s_intersect']:.1f} Mrays/s, gather {R['gathers_per_s']/1e3:.1f} k/s, build {R['kd_build_ms']/1e3:.1f} s) |
| step, 1 × B200: `e2e` (host buffers) | **{d['e2e']['value']:.0f} Mrays/s** ({d['e2e']['ms_per_step']:.0f} ms; compact records: {d['e2e']['h2d_bytes_per_step']/1e9:.2f} GB in + {d['e2e']['d2h_bytes_per_step']/1e9:.2f} GB out per step); with the reference-shaped 72 / 136-byte records {d['e2e']['reference_shaped_records']['value']:.0f} Mrays/s ({d['e2e']['reference_shaped_records']['h2d_bytes_per_step']/1e9:.2f} + {d['e2e']['reference_shaped_records']['d2h_bytes_per_step']/1e9:.2f} GB) | same (host memory is where it lives) |
| closest hit, jensen primary rays | {d['mrays_per_s_intersect']/1e3:.1f} Grays/s | {R['mrays_per_s_intersect']:.1f} Mrays/s |
| gather, k 225 on 10⁷ photons | {d['gathers_per_s']/1e6:.0f} M gathers/s = {d['roofline']['achieved']:.0f} GB/s algorithmic = {100*d['roofline']['frac']:.1f} % of the measured 6543 GB/s | {R['gathers_per_s']/1e3:.0f} k/s |
| gather, k 50 on 2 176 photons (2²³ queries) | {d['gathers_small_map']['gathers_per_s']/1e6:.0f} M gathers/s (443 M with the tree walk); {d['gathers_small_map']['roofline']['warp_instructions_per_gather']:.0f} warp-instructions per gather = {100*d['gathers_small_map']['roofline']['frac']:.0f} % of the issue rate | — |
| closest hit, violin.scn (104 300 triangles), 2048² | camera rays {d['intersect_mesh']['camera']['mrays_per_s']:.0f} Mrays/s, random rays {d['intersect_mesh']['random']['mrays_per_s']:.0f} Mrays/s (round-1 code: 1951 / 983) | ≈ 3.8 k rays/s (brute force over the mesh, SURVEY §3) |
| kd-tree build, 10⁷ photons | {d['stage_ms']['kd_build']:.1f} ms (round 1: 41.8) | {R['kd_build_ms']/1e3:.1f} s |
| frame: cornell 512² aa 1, all flags default (configs[0]) | {f5['frame_ms']:.0f} ms = photon pass {f5['photon_pass_ms']:.0f} + render {f5['render_ms']:.0f} (round 1: 1316 = 239 + 1077) | ≈ 183 s + ≈ 2140 s extrapolated from BASELINE.md's 128² run |
| **the metric's frame: cornell 1024² aa 2, all flags default** | **{f1['frame_ms']/1e3:.2f} s** on 1 GPU (round 1: 17.0 s), {f8['cornell_1024_aa2']['frame_ms']/1e3:.2f} s on 8 | days |
"""
rows = ["| GPUs | step `value` (Mrays/s) | weak eff. | `e2e`, compact records | `e2e`, reference-shaped records | cornell 512² aa 1 frame | cornell 1024² aa 2 frame | strong eff. of that frame |", "|---|---|---|---|---|---|---|---|"]
for n in (1, 2, 4, 8):
    b = B[n]; fr = b["frames"]
    rows.append(f"| {n} | {b['value']:.0f} | {b['value']/(n*B[1]['value']):.3f} | {b['e2e']['value']:.0f} | {b['e2e']['reference_shaped_records']['value']:.0f} | {fr['cornell_512_aa1']['frame_ms']:.0f} ms | {fr['cornell_1024_aa2']['frame_ms']/1e3:.2f} s | {B[1]['frames']['cornell_1024_aa2']['frame_ms']/(n*fr['cornell_1024_aa2']['frame_ms']):.2f} |")
dl = f8["distributed_lit_4096_aa2_10x_photons"]
scaling = ("Numbers (`profiles/r02_bench_n{1,2,4,8}.json`):\n\n" + "\n".join(rows) + "\n\n`value` scales because nothing but the replicated kd build and a 320 MB all-gather (1 ms over NVLink) is added per rank; `e2e` is bounded "
  "by the host at N = 8 (eight processes move gigabytes each per step through one host's memory system: the compact records lift it from "
  f"{B[8]['e2e']['reference_shaped_records']['value']:.0f} to {B[8]['e2e']['value']:.0f} Mrays/s); the frame's photon pass "
  f"goes {B[1]['frames']['cornell_1024_aa2']['photon_pass_ms']:.0f} → {B[8]['frames']['cornell_1024_aa2']['photon_pass_ms']:.0f} ms (the kd build and the emission rounds' fixed costs do not shrink).  `image_mean` is identical at every N for every frame.\n\n"
  "BASELINE.json configs 3–5 as whole-program frames, every flag at its default (`bench.py --frames all`; 1 GPU / 8 GPUs): "
  f"jensen 1024² aa 2 {d['frames']['jensen_1024_aa2']['frame_ms']/1e3:.2f} s / {f8['jensen_1024_aa2']['frame_ms']/1e3:.2f} s; teapot 1024² {d['frames']['teapot_1024']['frame_ms']:.0f} ms / {f8['teapot_1024']['frame_ms']:.0f} ms; "
  f"stilllife 2048² {d['frames']['stilllife_2048']['frame_ms']/1e3:.2f} s / {f8['stilllife_2048']['frame_ms']/1e3:.2f} s; violin 2048² {d['frames']['violin_2048']['frame_ms']/1e3:.2f} s / {f8['violin_2048']['frame_ms']/1e3:.2f} s; "
  f"the lit variant of distributed.scn at 4096² aa 2 with `-caustic 100000000` (config 5, 8 GPUs only): {dl['frame_ms']/1e3:.1f} s "
  f"(photon pass {dl['photon_pass_ms']/1e3:.2f} s for {dl['photons_stored']['caustic']} caustic photons, render {dl['render_ms']/1e3:.1f} s).  "
  "Stock `distributed.scn` has a black light: the reference segfaults on it with default flags (SURVEY §0), here it exits cleanly with the photon stages off (`tests/test_gpu_cli.py`).\n\n")
s = open(P("DESIGN.md")).read()
a = s.index("| | this repo | reference, 16 host cores |"); b = s.index("Per-stage device times of the metric's frame")
s = s[:a] + table + "\n" + s[b:]
a = s.index("Per-stage device times of the metric's frame"); b = s.index("The target of the north star (100 ms on one B200)")
s = s[:a] + "Per-stage device times of the metric's frame (1 GPU, `frames.cornell_1024_aa2.stage_ms_max_over_ranks`; the finishing pass overlaps the gather estimates): " + stages + ".\n" + s[b:]
a = s.index("Numbers (`profiles/r02_bench_n{1,2,4,8}.json`):"); b = s.index("## 7. The widened rows")
s = s[:a] + scaling + s[b:]
open(P("DESIGN.md"), "w").write(s)
r = open(P("README.md")).read()
a = r.index("Numbers of the last capture")
r = r[:a] + ("Numbers of the last capture (B200, `profiles/r02_*`, details in `DESIGN.md` §5): "
  f"one step (photon-map replicate + kd build + closest hit + gather for a 1024² aa 2 jensen frame) runs at {d['value']:.0f} Mrays/s with resident inputs "
  f"(kd build of 10⁷ photons {d['stage_ms']['kd_build']:.0f} ms, closest hit {d['mrays_per_s_intersect']/1e3:.1f} Grays/s, gather {d['gathers_per_s']/1e6:.0f} M gathers/s at k = 225 = {100*d['roofline']['frac']:.0f} % of the measured HBM roofline, issue-bound) and {d['e2e']['value']:.0f} Mrays/s end to end through host "
  f"buffers, against {R['value']:.3f} Mrays/s for the reference's own code on the box's 16 host cores; 8 GPUs: {B[8]['value']:.0f} Mrays/s.  The metric's frame — cornell 1024² aa 2 with every flag at its default — renders in "
  f"{f1['frame_ms']/1e3:.2f} s on one B200 (17.0 s at the end of round 1) and {f8['cornell_1024_aa2']['frame_ms']/1e3:.2f} s on eight; cornell 512² aa 1: {f5['photon_pass_ms']/1e3:.2f} s photon pass + {f5['render_ms']/1e3:.2f} s render.\n")
open(P("README.md"), "w").write(r)
print("filled: value", round(d["value"], 1), "frame", round(f1["frame_ms"]), "ms")
