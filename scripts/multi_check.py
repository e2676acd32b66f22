"""Launched by torch.distributed.run with N ranks (one per GPU): photon pass + render sharded by the C ABI over NCCL
(gi_comm_init_rank; csrc/comm.cu), checked against the same work done by ONE context without a communicator:
  * the photon maps are byte-identical for any N (photon t of a round goes to rank t mod N, counts all-reduced per round);
  * every rank holds the same maps;
  * the frame gathered on rank 0 equals the one-GPU frame (<= 1 LSB where FP64 atomic-add order differs).
Exit code 0 = all checks passed."""
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
    multi_gpu.attach_communicator(ctx, dist, torch, dev)
    assert ctx.comm_info() == (rank, world)
    flags = "-global 20000 -caustic 200000 -it 4 -st 2 -tt 2 -lt 4 -ss 4 -md 4 -seed 5".split()
    params, _ = gi_b200.parse_flags(flags)
    st = multi_gpu.trace_photons_sharded(ctx, params)
    assert st.global_stored >= 20000 and st.caustic_stored >= 200000
    maps = [ctx.get_photons(0), ctx.get_photons(1)]
    assert len(maps[0]) == st.global_stored and len(maps[1]) == st.caustic_stored
    # every rank holds the same replicated maps
    h = torch.tensor([int(np.frombuffer(m.tobytes(), dtype=np.uint8).astype(np.int64).sum()) for m in maps], device=dev)
    hs_ = [torch.zeros_like(h) for _ in range(world)]
    dist.all_gather(hs_, h)
    assert all(bool((x == hs_[0]).all()) for x in hs_), "photon maps differ between ranks"
    img, rs = multi_gpu.render_sharded(ctx, params, 48, 32, 1)
    if rank == 0:
        # the same job on ONE context without a communicator
        solo = gi_b200.Context(local)
        solo.upload_scene(hs)
        p1, _ = gi_b200.parse_flags(flags)
        s1 = solo.trace_photons(p1)
        assert (s1.global_stored, s1.caustic_stored, s1.global_emitted, s1.caustic_emitted, s1.rounds) == \
               (st.global_stored, st.caustic_stored, st.global_emitted, st.caustic_emitted, st.rounds), "the adaptive schedule depends on N"
        for k in range(2):
            assert solo.get_photons(k).tobytes() == maps[k].tobytes(), "the photon map depends on the number of ranks"
        p1.indirect_illum, p1.caustic_illum, p1.direct_photon_illum = s1.indirect_illum, s1.caustic_illum, s1.direct_photon_illum
        one, r1 = solo.render(p1, 48, 32, 1)
        assert img.shape == one.shape == (32, 48, 3)
        d = np.abs(img.astype(int) - one.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 0.01, (d.max(), (d > 0).mean())
        for f in ("screen_rays", "shadow_rays", "monte_rays", "transmissive_samples", "specular_samples", "indirect_samples", "caustic_samples"):
            assert getattr(rs, f) == getattr(r1, f), f
        solo.close()
        print("multi_check OK: world", world, "photons", st.global_stored, st.caustic_stored, "mean", img.mean())
    # gi_load_photons_sharded_device: ragged shares (rank r holds 1000 + 317 r photons) are concatenated in rank order on every rank
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from gi_testlib import synthetic_photons
    shares = [synthetic_photons(1000 + 317 * r, 40 + r) for r in range(world)]
    mine = torch.from_numpy(shares[rank].view(np.uint8).reshape(-1, 32).copy()).to(dev)
    ctx.load_photons_sharded_device(gi_b200.GI_MAP_CAUSTIC, mine.data_ptr(), len(shares[rank]))
    got = ctx.get_photons(gi_b200.GI_MAP_CAUSTIC)
    assert got.tobytes() == np.concatenate(shares).tobytes(), "sharded load: not the shares in rank order"
    ctx.build_maps()
    if rank == 0:
        print("sharded load OK:", len(got), "photons")
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
