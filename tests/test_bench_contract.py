"""bench.py's reference arm on the CPU (the driver runs `bench.py --impl reference` beside ours): with tiny sizes it must
print ONE JSON line carrying the contract's keys, a cpu_baseline that describes itself, an e2e object with zero copy bytes,
and unit counts formed the way the GPU arm forms them (a gather at every DIFFUSE primary hit).  Without the compiled
reference (oracle/_ref) it must say so and exit 0."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-rays", "40000",
                        "--cpu-queries", "1500", "--photons", "60000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    if "unavailable" in d:
        assert not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_harness"))
        return
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
                "config", "cpu_baseline", "e2e", "step_units"):
        assert key in d, key
    assert d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["vs_baseline"] is None and d["dtype"] == "f64" and d["value"] > 0
    assert "workload" in d["config"] and "jensen" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "stock input/jensen.scn" in cb["sample"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    u = d["step_units"]
    # jensen 1024 x 1024 aa 2: every primary ray, a gather at every diffuse hit (the two spheres and the misses take none), one build
    assert u["rays"] == 4096 * 4096 and u["map_builds"] == 1 and 0.85 * u["rays"] < u["gathers"] < 0.95 * u["rays"]
    # value = units / (rays / isect rate + gathers / gather rate + build)
    sec = u["rays"] / (d["mrays_per_s_intersect"] * 1e6) + u["gathers"] / d["gathers_per_s"] + d["kd_build_ms"] * 1e-3
    assert abs(d["value"] - (u["rays"] + u["gathers"]) / sec / 1e6) <= 1e-9 * d["value"]
