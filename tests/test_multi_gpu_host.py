"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in global-illumination_b200/multi_gpu.py: the ragged
photon all-gather, the emitted-count reduction and power scale, and the frame assembly -- with a stand-in context
that records what the C ABI would be asked to do (no kernels run here; the GPU path is tests/test_gpu_multi.py)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeStats:
    def __init__(self, rank):
        self.global_emitted = 1000 + 10 * rank; self.caustic_emitted = 0 if rank else 0
        self.total_power = 12.5; self.trace_ms = 1.0
        self.global_stored = 0; self.caustic_stored = 0


class FakeContext:
    """photons live in numpy arrays; `ptr` arguments are torch tensors' data_ptr() of CPU tensors"""
    def __init__(self, rank, torch):
        self.rank, self.torch = rank, torch
        self.maps = {0: None, 1: None}; self.scaled = {}; self.built = False; self.render_calls = []
        self._tensors = {}

    def trace_photons(self, params):
        assert (params.rank, params.world_size) == (self.rank, 2)
        rng = np.random.default_rng(self.rank)
        self.maps[0] = rng.integers(0, 255, size=(5 + 3 * self.rank, 32), dtype=np.uint8)      # ragged: 5 and 8 photons
        self.maps[1] = np.zeros((0, 32), dtype=np.uint8)                                       # no caustic photon anywhere
        return FakeStats(self.rank)

    def photon_count(self, which):
        return len(self.maps[which])

    def get_photons_device(self, which, ptr, capacity, stream=None):
        import ctypes
        ctypes.memmove(ptr, self.maps[which].ctypes.data, self.maps[which].nbytes)

    def load_photons_device(self, which, ptr, n):
        import ctypes
        buf = np.zeros((n, 32), dtype=np.uint8)
        if n:
            ctypes.memmove(buf.ctypes.data, ptr, buf.nbytes)
        self.maps[which] = buf

    def scale_photon_power(self, which, factor):
        self.scaled[which] = factor

    def build_maps(self, params=None):
        self.built = True

    def render(self, params, width, height, aa):
        import multi_gpu
        img = np.zeros((height, width, 3), dtype=np.uint8)
        img[:, list(multi_gpu.columns_of(width, params.rank, params.world_size))] = 50 + 100 * params.rank
        self.render_calls.append((params.rank, params.world_size))
        return img, None


def _worker(rank, port, out):
    sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
    import torch
    import torch.distributed as dist
    import gi_b200
    import multi_gpu
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE="2")
    dist.init_process_group("gloo", rank=rank, world_size=2)
    dev = torch.device("cpu")
    res = {}
    # ragged gather
    local = torch.arange((3 + 2 * rank) * 4, dtype=torch.int64).reshape(-1, 4) + 1000 * rank
    full, counts = multi_gpu.gather_ragged(dist, torch, local)
    res["counts"] = counts; res["full"] = full.numpy().copy()
    empty, c2 = multi_gpu.gather_ragged(dist, torch, torch.zeros((0, 4), dtype=torch.int64))
    res["empty"] = (tuple(empty.shape), c2)
    # photon pass
    ctx = FakeContext(rank, torch)
    params = gi_b200.GiParams(); params.indirect_illum = 1; params.caustic_illum = 1
    totals = multi_gpu.trace_photons_sharded(ctx, params, dist, torch, dev)
    res["totals"] = totals; res["global_map"] = ctx.maps[0].copy(); res["scaled"] = dict(ctx.scaled); res["built"] = ctx.built
    res["flags"] = (params.indirect_illum, params.caustic_illum, params.rank, params.world_size)
    # frame assembly
    img, _ = multi_gpu.render_sharded(ctx, params, 7, 3, 1, dist, torch, dev)
    res["img"] = img
    out.put((rank, res))
    dist.destroy_process_group()


def test_world_size_2_host_logic():
    import torch.multiprocessing as mp
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctxm.Process(target=_worker, args=(r, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in range(2):
        res = got[r]
        assert res["counts"] == [3, 5]
        want = np.concatenate([np.arange(12).reshape(3, 4), np.arange(20).reshape(5, 4) + 1000])
        assert np.array_equal(res["full"], want)
        assert res["empty"] == ((0, 4), [0, 0])
        t = res["totals"]
        assert t["global_emitted"] == 2010 and t["global_stored"] == 13 and t["global_per_rank"] == [5, 8]
        assert t["caustic_stored"] == 0
        assert res["scaled"] == {0: pytest.approx(12.5 / 2010)}          # total_power / emitted summed over ranks
        assert res["built"] and res["flags"] == (1, 0, 0, 1)             # empty caustic map switches caustics off
    # both ranks hold the same replicated map: rank 0's photons then rank 1's
    assert np.array_equal(got[0]["global_map"], got[1]["global_map"])
    rng0, rng1 = np.random.default_rng(0), np.random.default_rng(1)
    assert np.array_equal(got[0]["global_map"][:5], rng0.integers(0, 255, size=(5, 32), dtype=np.uint8))
    assert np.array_equal(got[0]["global_map"][5:], rng1.integers(0, 255, size=(8, 32), dtype=np.uint8))
    img = got[0]["img"]
    assert got[1]["img"] is None and img.shape == (3, 7, 3)
    assert np.all(img[:, 0::2] == 50) and np.all(img[:, 1::2] == 150)


def test_column_ranges_partition_the_frame():
    sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
    import multi_gpu
    for width in (1, 7, 64, 1023):
        for world in (1, 2, 3, 8):
            cols = []
            for r in range(world):
                cols += list(multi_gpu.columns_of(width, r, world))
            assert sorted(cols) == list(range(width))
