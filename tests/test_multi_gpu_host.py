"""world_size-2 gloo tests (CPU) of the multi-GPU host side (global-illumination_b200/multi_gpu.py): the communicator
bootstrap over torch.distributed, and the two partitions the C side applies -- photons t = rank (mod world) of every
emission round, pixel columns x = rank (mod world) -- replayed with a stand-in context whose "collectives" are gloo
calls.  No kernels run here; the GPU path is tests/test_gpu_multi.py (NCCL)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeContext:
    """records what the C ABI would be asked to do; gi_trace_photons / gi_render replayed on the CPU with gloo standing in
    for the NCCL calls of csrc/comm.cu (same partitions, same assembly)"""
    def __init__(self, dist, torch):
        self.dist, self.torch = dist, torch
        self.comm = None; self.map = None

    def comm_init_rank(self, unique_id, rank, world):
        assert len(unique_id) == 128
        self.comm = (bytes(unique_id), rank, world)

    def comm_info(self):
        return self.comm[1], self.comm[2]

    def trace_photons(self, params):
        """rounds of n photons; photon t 'stores' iff hash(t) is small; per-round all-reduce of the counts; final gather +
        sort by photon number -- trace_map of csrc/photon_trace.cu in miniature"""
        import multi_gpu
        torch, dist = self.torch, self.dist
        _, rank, world = self.comm
        target, stored, emitted, mine = 40, 0, 0, []
        rate = 4.0
        rounds = 0
        while stored < target:
            n = int((target - stored) / rate + 1) * 8
            ids = [emitted + t for t in multi_gpu.round_share(n, rank, world)]
            mine += [i for i in ids if (i * 2654435761) % 7 == 0]
            c = torch.tensor([len(mine)], dtype=torch.int64); dist.all_reduce(c)
            stored = int(c.item()); emitted += n; rounds += 1
            rate = max(stored / emitted, 1e-3)
        counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([len(mine)], dtype=torch.int64))
        counts = [int(x.item()) for x in counts]
        pad = torch.zeros(max(counts), dtype=torch.int64); pad[:len(mine)] = torch.tensor(mine, dtype=torch.int64)
        out = [torch.zeros(max(counts), dtype=torch.int64) for _ in range(world)]
        dist.all_gather(out, pad)
        self.map = np.sort(np.concatenate([o.numpy()[:c] for o, c in zip(out, counts)]))

        class St:
            pass
        st = St(); st.total_power = 12.5; st.indirect_illum, st.caustic_illum, st.direct_photon_illum = 1, 0, 0
        st.global_stored, st.global_emitted, st.rounds = stored, emitted, rounds
        return st

    def render(self, params, width, height, aa, want_image=True):
        import multi_gpu
        torch, dist = self.torch, self.dist
        _, rank, world = self.comm
        cols = list(multi_gpu.columns_of(width, rank, world))
        strip = np.zeros((height, len(cols), 3), dtype=np.uint8)
        for l, x in enumerate(cols):
            strip[:, l] = (x * 7 + np.arange(height)[:, None] * 3 + np.arange(3)) % 251
        ncols = [len(list(multi_gpu.columns_of(width, r, world))) for r in range(world)]
        out = [torch.zeros(height * max(ncols) * 3, dtype=torch.uint8) for _ in range(world)]
        pad = torch.zeros(height * max(ncols) * 3, dtype=torch.uint8); pad[:strip.size] = torch.from_numpy(strip.reshape(-1))
        dist.all_gather(out, pad)
        if not want_image:
            return None, None
        img = np.zeros((height, width, 3), dtype=np.uint8)
        for r in range(world):
            s = out[r].numpy()[: height * ncols[r] * 3].reshape(height, ncols[r], 3)
            img[:, list(multi_gpu.columns_of(width, r, world))] = s
        return img, None


def _worker(rank, world, port, out):
    sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
    import torch
    import torch.distributed as dist
    import gi_b200
    import multi_gpu
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cpu")
    res = {}
    ctx = FakeContext(dist, torch)
    # bootstrap: rank 0's id reaches everyone (the real gi_comm_unique_id needs NCCL; any 128 bytes do here)
    orig = gi_b200.comm_unique_id
    gi_b200.comm_unique_id = lambda: bytes((7 * i + 3) % 256 for i in range(128))
    try:
        res["rank_world"] = multi_gpu.attach_communicator(ctx, dist, torch, dev)
    finally:
        gi_b200.comm_unique_id = orig
    res["comm"] = ctx.comm
    params = gi_b200.GiParams(); params.indirect_illum = 1; params.caustic_illum = 1
    st = multi_gpu.trace_photons_sharded(ctx, params)
    res["map"] = ctx.map; res["flags"] = (params.indirect_illum, params.caustic_illum); res["emitted"] = st.global_emitted
    res["rounds"] = st.rounds
    img, _ = multi_gpu.render_sharded(ctx, params, 7, 3, 1)
    res["img"] = img
    out.put((rank, res))
    dist.destroy_process_group()


def _run(world):
    import torch.multiprocessing as mp
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29500 + (os.getpid() * 7 + world) % 2000
    procs = [ctxm.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return got


def test_world_size_2_host_logic():
    got = _run(2)
    want_id = bytes((7 * i + 3) % 256 for i in range(128))
    for r in range(2):
        assert got[r]["rank_world"] == (r, 2)
        assert got[r]["comm"] == (want_id, r, 2)
        assert got[r]["flags"] == (1, 0)                   # an empty caustic map switches caustics off on every rank
    assert np.array_equal(got[0]["map"], got[1]["map"])    # replicated
    img = got[0]["img"]
    assert got[1]["img"] is None and img.shape == (3, 7, 3)
    want = (np.arange(7)[None, :, None] * 7 + np.arange(3)[:, None, None] * 3 + np.arange(3)[None, None, :]) % 251
    assert np.array_equal(img, want.astype(np.uint8))
    # the same schedule and the same map for any number of ranks: replay with one rank in this process
    one = _run(1)
    assert np.array_equal(one[0]["map"], got[0]["map"]) and one[0]["emitted"] == got[0]["emitted"] and one[0]["rounds"] == got[0]["rounds"]
    assert np.array_equal(one[0]["img"], img)


def test_partitions_cover_everything_once():
    sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
    import multi_gpu
    for n in (0, 1, 7, 64, 1023):
        for world in (1, 2, 3, 8):
            cols, ph = [], []
            for r in range(world):
                cols += list(multi_gpu.columns_of(n, r, world)); ph += list(multi_gpu.round_share(n, r, world))
                # the counts the C side computes in closed form (render.cu: local columns, photon_trace.cu: n_local)
                assert len(list(multi_gpu.columns_of(n, r, world))) == (n - r + world - 1) // world if n > r else True
            assert sorted(cols) == list(range(n)) and sorted(ph) == list(range(n))
