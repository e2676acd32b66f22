"""One process per GPU: the two places where the renderer shards.

  photon pass   every rank traces its share of the photon budget (gi_trace_photons with rank / world_size, the
                reference's per-thread split, photonmap.cpp:289-321); ONE all-gather per map replicates the photons
                (ragged: ranks store different counts), the emitted counts are all-reduced, every rank applies the
                power scale and builds its own kd-trees.  No collective touches the gather / render data path.
  render        rank r renders the pixel columns x = r (mod world) (gi_render with rank / world_size) into a
                zero-initialised frame; the frames are summed onto rank 0.

torch.distributed is the plumbing (NCCL on the GPUs; the same functions run under gloo on CPU tensors for the tests).
"""
import numpy as np

import gi_b200


def gather_ragged(dist, torch, local):
    """all-gather of per-rank row blocks with different row counts: returns (rows of every rank concatenated in rank
    order, list of per-rank counts).  `local` is a 2-D tensor on the backend's device."""
    world = dist.get_world_size()
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    m = max(counts + [1])
    padded = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    out = torch.empty((world * m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded)
    parts = [out[r * m: r * m + counts[r]] for r in range(world)]
    return torch.cat(parts, dim=0).contiguous(), counts


def columns_of(width, rank, world):
    """output pixel columns of rank `rank` -- the split gi_render applies (csrc/render.cu): x = rank (mod world), the
    reference's own interleaving of columns over threads (render.cpp:90).  Whole pixel columns, so every 2^aa x 2^aa
    box-filter sum stays on one rank and the ranks' frames are disjoint."""
    return range(rank, width, world)


def trace_photons_sharded(ctx, params, dist, torch, device):
    """MapPhotons across the ranks of the default process group.  Returns a dict of totals; the maps are left built on
    every rank and `params`' illumination flags are updated like photonmap.cpp:339-361."""
    world, rank = dist.get_world_size(), dist.get_rank()
    params.rank, params.world_size = rank, world
    st = ctx.trace_photons(params)
    emitted = torch.tensor([st.global_emitted, st.caustic_emitted], dtype=torch.int64, device=device)
    dist.all_reduce(emitted)
    totals = {"global_emitted": int(emitted[0].item()), "caustic_emitted": int(emitted[1].item()), "total_power": st.total_power,
              "local_trace_ms": st.trace_ms}
    stream = torch.cuda.current_stream().cuda_stream if device.type == "cuda" else None
    for which, name in ((gi_b200.GI_MAP_GLOBAL, "global"), (gi_b200.GI_MAP_CAUSTIC, "caustic")):
        n = ctx.photon_count(which)
        local = torch.empty((n, 32), dtype=torch.uint8, device=device)
        if n:
            ctx.get_photons_device(which, local.data_ptr(), n, stream)
        full, counts = gather_ragged(dist, torch, local)
        torch.cuda.synchronize(device) if device.type == "cuda" else None
        ctx.load_photons_device(which, full.data_ptr(), full.shape[0])
        e = totals[name + "_emitted"]
        if full.shape[0] and e:
            ctx.scale_photon_power(which, st.total_power / e)
        totals[name + "_stored"] = int(full.shape[0]); totals[name + "_per_rank"] = counts
    if st.total_power > 0:
        wanted_global = params.indirect_illum or params.direct_photon_illum
        if wanted_global and totals["global_stored"] == 0:
            params.indirect_illum = 0; params.direct_photon_illum = 0
        if params.caustic_illum and totals["caustic_stored"] == 0:
            params.caustic_illum = 0
    ctx.build_maps(params)
    params.rank, params.world_size = 0, 1
    return totals


def render_sharded(ctx, params, width, height, aa, dist, torch, device):
    """RenderImage across the ranks: returns (uint8 frame [H, W, 3] on rank 0, None elsewhere; GiRenderStats of this rank)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    params.rank, params.world_size = rank, world
    img, st = ctx.render(params, width, height, aa)
    params.rank, params.world_size = 0, 1
    # the ranks' column ranges are disjoint and everything outside a rank's range is zero: a sum assembles the frame
    t = torch.from_numpy(img).to(device).to(torch.int32)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank != 0:
        return None, st
    return t.to(torch.uint8).cpu().numpy(), st
