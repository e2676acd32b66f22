import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
ctx = gi_b200.Context(0)
for scene in sys.argv[1:] or ["cornell", "jensen"]:
    hs = gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", scene + ".gisc"))
    ctx.upload_scene(hs)
    params, _ = gi_b200.parse_flags([])
    for rep in range(2):
        t0 = time.perf_counter(); st = ctx.trace_photons(params); t1 = time.perf_counter()
    print(f"{scene}: stored g {st.global_stored} c {st.caustic_stored} emitted g {st.global_emitted} c {st.caustic_emitted} rounds {st.rounds} "
          f"trace {st.trace_ms:.1f} ms build {st.build_ms:.1f} ms wall {1e3*(t1-t0):.1f} ms -> {st.caustic_emitted/st.trace_ms/1e3:.1f} M photons/s emitted")
