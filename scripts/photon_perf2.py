"""photon pass timing per scene with stage times (second of two passes)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
from gi_testlib import golden_gisc
for scene in sys.argv[1:] or ["cornell", "jensen", "stilllife"]:
    ctx = gi_b200.Context(0)
    ctx.upload_scene(gi_b200.HostScene(golden_gisc(scene)))
    ctx.stage_timing(True)
    for rep in range(2):
        params, _ = gi_b200.parse_flags(["-seed", "1"])
        t0 = time.perf_counter(); st = ctx.trace_photons(params); t1 = time.perf_counter()
    s = ctx.stage_times()
    print(scene, "photon pass %.1f ms (trace %.1f emit_filter %.1f kd %.1f) stored %d %d" % ((t1 - t0) * 1e3, s["trace_ms"], s["emit_filter_ms"], s["kd_build_ms"], st.global_stored, st.caustic_stored))
    ctx.close()
