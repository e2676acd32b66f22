"""N = 2 on real GPUs (skipped on a one-GPU box): the multi-GPU split of the C ABI (csrc/comm.cu) over NCCL, both ways
of forming the communicator.  The host logic is covered on CPU by tests/test_multi_gpu_host.py (gloo)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PHOTONMAP = os.path.join(ROOT, "global-illumination_b200", "bin", "photonmap")


def _gpus():
    import torch
    return torch.cuda.device_count()


def test_two_rank_job_equals_one_gpu_job():
    """one process per GPU (torchrun), gi_comm_init_rank: maps byte-identical to the one-GPU maps, frame equal"""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29731", os.path.join(ROOT, "scripts", "multi_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "multi_check OK" in r.stdout and "sharded load OK" in r.stdout


def test_cli_gpus_flag_renders_the_same_image(tmp_path):
    """one process driving N devices (`photonmap -gpus 2`: gi_comm_init_all + a host thread per GPU) writes the image
    the one-GPU program writes, and the same photon map file"""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from gi_testlib import GOLDEN, read_png_rgb
    scn = os.path.join(GOLDEN, "scn", "cornell.scn")
    common = "-resolution 40 40 -aa 1 -global 3000 -caustic 50000 -it 4 -st 2 -tt 2 -lt 4 -ss 4 -md 4 -seed 9 -v".split()
    outs = []
    for n in (1, 2):
        png, gp = tmp_path / f"o{n}.png", tmp_path / f"c{n}.gipm"
        r = subprocess.run([PHOTONMAP, scn, str(png), *common, "-gpus", str(n), "-save_caustic", str(gp)], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        outs.append((read_png_rgb(png), open(gp, "rb").read(), [l for l in r.stdout.splitlines() if (l.strip().startswith("#") or "Total" in l) and "Time" not in l]))
    assert outs[0][1] == outs[1][1], "the caustic map depends on the number of GPUs"
    assert [l for l in outs[0][2] if "GPUs" not in l] == [l for l in outs[1][2] if "GPUs" not in l]      # the -v counters
    d = np.abs(outs[0][0].astype(int) - outs[1][0].astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01


def test_comm_errors(gpu_ctx):
    import gi_b200
    with pytest.raises(gi_b200.GiError, match="two contexts on one device"):
        gi_b200.comm_init_all([gpu_ctx, gpu_ctx])
    assert gpu_ctx.comm_info() == (0, 1)
