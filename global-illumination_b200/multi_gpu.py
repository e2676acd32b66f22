"""One process per GPU (torchrun): the host side of the multi-GPU split.

The split itself lives in the C ABI (csrc/comm.cu, gi_b200.h "Multi-GPU"): with an NCCL communicator attached to the
context, gi_trace_photons shards every emission round over the ranks (all-reduce of the stored counts per round, one
ragged all-gather of the photons at the end, kd-trees built on every rank) and gi_render renders the pixel columns
x = rank (mod world) and gathers the 8-bit strips to rank 0.  What is left for the host is the bootstrap: rank 0 makes
the ncclUniqueId and the launcher's own channel carries its 128 bytes to the other ranks -- torch.distributed here
(NCCL on the GPUs; gloo on CPU tensors in the tests), MPI or a file elsewhere.
"""
import numpy as np

import gi_b200


def columns_of(width, rank, world):
    """output pixel columns of rank `rank` -- the split gi_render applies (csrc/render.cu): x = rank (mod world), the
    reference's own interleaving of columns over threads (render.cpp:90).  Whole pixel columns, so every 2^aa x 2^aa
    box-filter sum stays on one rank and the ranks' strips are disjoint."""
    return range(rank, width, world)


def round_share(n, rank, world):
    """photons of an emission round of n photons that rank `rank` traces: t = rank (mod world), t < n
    (csrc/photon_trace.cu: trace_map)"""
    return range(rank, n, world)


def attach_communicator(ctx, dist, torch, device):
    """Join this rank's context to the job-wide NCCL communicator: rank 0 creates the unique id, torch.distributed
    broadcasts it, every rank calls gi_comm_init_rank.  Returns (rank, world)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    buf = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        buf.copy_(torch.from_numpy(np.frombuffer(gi_b200.comm_unique_id(), dtype=np.uint8).copy()))
    dist.broadcast(buf, src=0)
    ctx.comm_init_rank(bytes(buf.cpu().numpy().tobytes()), rank, world)
    return rank, world


def trace_photons_sharded(ctx, params):
    """MapPhotons across the ranks of the context's communicator.  Returns GiPhotonStats (global totals, the same on
    every rank); `params`' illumination flags are updated like photonmap.cpp:339-361."""
    st = ctx.trace_photons(params)
    if st.total_power > 0:
        params.indirect_illum, params.caustic_illum, params.direct_photon_illum = st.indirect_illum, st.caustic_illum, st.direct_photon_illum
    return st


def render_sharded(ctx, params, width, height, aa):
    """RenderImage across the ranks: (uint8 frame [H, W, 3] on rank 0, None elsewhere; GiRenderStats with the ray
    counters summed over the ranks)."""
    rank, _ = ctx.comm_info()
    return ctx.render(params, width, height, aa, want_image=(rank == 0))
