"""Render timing in one process (second run reported, so module loading and allocator warm-up are excluded).
usage: render_perf.py [scene] [res] [aa] [flags...]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
scene = sys.argv[1] if len(sys.argv) > 1 else "cornell"
res = int(sys.argv[2]) if len(sys.argv) > 2 else 256
aa = int(sys.argv[3]) if len(sys.argv) > 3 else 1
flags = sys.argv[4:]
ctx = gi_b200.Context(0)
hs = gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", scene + ".gisc"))
ctx.upload_scene(hs)
params, _ = gi_b200.parse_flags(flags)
st = ctx.trace_photons(params)
params.indirect_illum, params.caustic_illum, params.direct_photon_illum = st.indirect_illum, st.caustic_illum, st.direct_photon_illum
for rep in range(int(os.environ.get("REPS", "3"))):
    t0 = time.perf_counter(); img, rs = ctx.render(params, res, res, aa); t1 = time.perf_counter()
    total = rs.screen_rays + rs.shadow_rays + rs.monte_rays + rs.transmissive_samples + rs.specular_samples + rs.indirect_samples + rs.caustic_samples
    print(f"{scene} {res}^2 aa{aa} rep {rep}: render {rs.render_ms:.1f} ms (wall {1e3*(t1-t0):.1f}) ray-ops {total} -> {total/rs.render_ms/1e3:.1f} M/s; "
          f"gathers {rs.indirect_samples + rs.caustic_samples}")
