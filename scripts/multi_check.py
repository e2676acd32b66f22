"""Launched by torch.distributed.run with N ranks (one per GPU): sharded photon pass + sharded render of a small frame,
checked on rank 0 against the same frame rendered by one GPU.  Exit code 0 = all checks passed."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "global-illumination_b200"))
import gi_b200      # noqa: E402
import multi_gpu    # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = gi_b200.Context(local)
    hs = gi_b200.HostScene(os.path.join(ROOT, "tests", "golden", "cornell.gisc"))
    ctx.upload_scene(hs)
    flags = "-global 20000 -caustic 20000 -it 4 -st 2 -tt 2 -lt 4 -ss 4 -md 4 -seed 5".split()
    params, _ = gi_b200.parse_flags(flags)
    totals = multi_gpu.trace_photons_sharded(ctx, params, dist, torch, dev)
    assert totals["global_stored"] >= 20000 and len(totals["global_per_rank"]) == world
    assert ctx.photon_count(0) == totals["global_stored"] and ctx.photon_count(1) == totals["caustic_stored"]
    # every rank holds the same replicated maps
    h = torch.tensor(np.frombuffer(ctx.get_photons(0).tobytes(), dtype=np.uint8).astype(np.int64).sum(), device=dev)
    hs_ = [torch.zeros_like(h) for _ in range(world)]
    dist.all_gather(hs_, h)
    assert all(int(x) == int(hs_[0]) for x in hs_), "photon maps differ between ranks"
    img, st = multi_gpu.render_sharded(ctx, params, 48, 32, 1, dist, torch, dev)
    if rank == 0:
        one, _ = ctx.render(params, 48, 32, 1)      # same maps, same seed, one GPU
        assert img.shape == one.shape == (32, 48, 3)
        d = np.abs(img.astype(int) - one.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 0.01, (d.max(), (d > 0).mean())
        print("multi_check OK: world", world, "photons", totals["global_stored"], totals["caustic_stored"], "mean", img.mean())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
