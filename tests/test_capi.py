"""CPU suite: the C-ABI shared library loads and exports every symbol include/gi_b200.h declares; host-side entry
points work; without a CUDA device the product fails loudly (no CPU fallback)."""
import ctypes
import gzip
import os
import re
import subprocess

import numpy as np
import pytest

import gi_b200
from gi_testlib import ROOT, SCENE_TOOL

GOLDEN = os.path.join(ROOT, "tests", "golden")


def header_functions():
    src = open(os.path.join(ROOT, "include", "gi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gi_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = gi_b200.lib()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"libgi_b200.so does not export {n}"
    assert sorted(gi_b200.EXPORTS) == names
    assert L.gi_abi_version() == 7


def test_default_params_match_reference_defaults():
    p = gi_b200.default_params()          # src/photonmap.cpp:41-106
    assert (p.global_photon_count, p.caustic_photon_count) == (2176, 10000000)
    assert (p.global_estimate_size, p.caustic_estimate_size) == (50, 225)
    assert (p.global_estimate_dist, p.caustic_estimate_dist) == (2.5, 0.225)
    assert (p.light_test, p.shadow_test, p.indirect_test) == (128, 128, 256)
    assert p.prob_absorb == 0.005 and p.max_monte_depth == 128 and p.ir_air == 1.0


def test_struct_sizes_match_header():
    assert ctypes.sizeof(gi_b200.GiCamera) == 16 * 8
    assert ctypes.sizeof(gi_b200.GiScene) == 6 * 4 + 6 * 8 + 16 * 8 + 16 * 8
    assert gi_b200.HIT.itemsize == 72 and gi_b200.PHOTON.itemsize == 32 and gi_b200.QUERY.itemsize == 136


def test_host_scene_roundtrip(tmp_path):
    s = gi_b200.HostScene(os.path.join(GOLDEN, "stilllife.gisc"))
    assert s.view.n_prims > 0 and s.view.n_nodes == 6
    out = str(tmp_path / "copy.gisc")
    s.save(out)
    assert open(out, "rb").read() == open(os.path.join(GOLDEN, "stilllife.gisc"), "rb").read()


def test_loader_errors_are_reported():
    with pytest.raises(gi_b200.GiError, match="Unable to open file"):
        gi_b200.HostScene("/nonexistent/scene.scn")


@pytest.mark.parametrize("name", ["cornell", "stilllife", "teapot", "transform", "cylinder", "circle", "art2", "distributed_lit"])
def test_flattened_scene_equals_reference_scene_graph(name, tmp_path):
    """text dump of our flattened scene == `ref_harness dump_scene` of the reference's R3Scene (committed fixture)"""
    out = str(tmp_path / "dump.txt")
    subprocess.run([SCENE_TOOL, "dump", os.path.join(GOLDEN, name + ".gisc"), out], check=True)
    ref = gzip.open(os.path.join(GOLDEN, name + ".dump.txt.gz"), "rb").read()
    assert open(out, "rb").read() == ref


def test_no_cpu_fallback_without_gpu():
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(gi_b200.GiError, match="no CUDA device"):
        gi_b200.Context(0)
