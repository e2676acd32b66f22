#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_photons.py tests/test_gpu_cli.py tests/test_gpu_render.py -x -q 2>&1 | tail -2
cat > /tmp/pp.py <<'PY'
import os, sys, time
ROOT = os.getcwd()
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200
for scene in ("cornell", "jensen"):
    ctx = gi_b200.Context(0)
    ctx.upload_scene(gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", scene + ".gisc")))
    ctx.stage_timing(True)
    for rep in range(2):
        params, _ = gi_b200.parse_flags(["-seed", "1"])
        t0 = time.perf_counter(); st = ctx.trace_photons(params); t1 = time.perf_counter()
    s = ctx.stage_times()
    print(scene, "photon pass %.1f ms (trace %.1f emit_filter %.1f kd %.1f) stored %d %d emitted %d" % ((t1 - t0) * 1e3, s["trace_ms"], s["emit_filter_ms"], s["kd_build_ms"], st.global_stored, st.caustic_stored, st.caustic_emitted))
    ctx.close()
PY
for v in 1 0; do GI_PHOTON_MEGAKERNEL=$v python /tmp/pp.py 2>&1 | sed "s/^/mega=$v /"; done
