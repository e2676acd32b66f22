"""gi_gather_batch_device_noalloc (include/gi_b200.h): the gather with every piece of scratch memory in a file-scope
static device array of csrc/gather.cu -- for callers that cannot call cudaMalloc;
it has no counterpart in the reference.  Checked through the C ABI against the existing entry point,
gi_gather_batch_device, on the same device buffers: identical kNN index sets and counts, radiance to 1e-9 relative (the
tolerance of tests/test_gpu_gather.py: only the summation order may differ), on TWO calls in a row (the static array
keeps the previous call's contents).  This file is named to be collected last."""
import numpy as np
import pytest

import gi_b200
from gi_testlib import synthetic_photons, GI_QUERY

pytestmark = pytest.mark.gpu


def _queries(n, seed):
    rng = np.random.default_rng(seed)
    ph = synthetic_photons(n, seed + 1000)
    q = np.zeros(n, dtype=GI_QUERY)
    q["p"] = ph["position"] + rng.normal(scale=1e-3, size=(n, 3))
    nn = rng.normal(size=(n, 3)); nn /= np.linalg.norm(nn, axis=1)[:, None]
    b = rng.normal(size=(n, 3)); b /= np.linalg.norm(b, axis=1)[:, None]
    q["n"] = nn; q["bounce"] = b; q["cos_theta"] = rng.uniform(-1, 1, n)
    q["kd"] = (0.5, 0.6, 0.7); q["ks"] = (0.2, 0.1, 0.3); q["shininess"] = 10.0
    return q


def _run(ctx, fn, torch, d_q, n, k, r, filt, params, want_idx):
    """one device-buffer gather through `fn`; outputs are poisoned first so that a call that writes nothing is seen"""
    dev = d_q.device
    idx = torch.full((n, k), -7, dtype=torch.int32, device=dev) if want_idx else None
    cnt = torch.full((n,), -7, dtype=torch.int32, device=dev)
    rgb = torch.full((n, 3), float("nan"), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    fn(gi_b200.GI_MAP_CAUSTIC, d_q.data_ptr(), n, k, r, filt, params, idx.data_ptr() if want_idx else None, cnt.data_ptr(),
       rgb.data_ptr(), stream)
    torch.cuda.synchronize()
    return (idx.cpu().numpy() if want_idx else None), cnt.cpu().numpy(), rgb.cpu().numpy()


def _same(ref, got, rtol=1e-9):
    if ref[0] is not None:
        assert np.array_equal(ref[0], got[0]), "kNN index sets differ"
    assert np.array_equal(ref[1], got[1]), "counts differ"
    scale = np.abs(ref[2]).max(axis=1, keepdims=True)
    assert np.isfinite(got[2]).all()
    assert (np.abs(got[2] - ref[2]) <= rtol * scale + 1e-300).all()


def test_noalloc_gather_equals_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    ctx = gi_b200.Context(0)                                # own context: the same whether run alone or after the suite
    try:
        ph = synthetic_photons(100000, seed=4242)
        params = gi_b200.default_params()
        q_all = _queries(6000, seed=9)
        # before any map is built on this context the direction table does not exist: refused, not allocated
        d_probe = torch.from_numpy(q_all[:8].view(np.uint8).reshape(8, 136).copy()).cuda()
        with pytest.raises(gi_b200.GiError, match="gi_build_maps"):
            _run(ctx, ctx.gather_device_noalloc, torch, d_probe, 8, 5, 0.1, 0, params, False)
        ctx.load_photons(gi_b200.GI_MAP_CAUSTIC, ph)
        ctx.build_maps()
        # 6000 queries: the Morton-ordered path (keys, permutation and sort storage in the static array);
        # 600: the unordered path (only the chunk counter)
        for nq, k, r, filt, want_idx in ((6000, 225, 0.2, 0, False), (6000, 50, 0.1, 2, True), (600, 225, 0.2, 1, True)):
            d_q = torch.from_numpy(q_all[:nq].view(np.uint8).reshape(nq, 136).copy()).cuda()
            ref = _run(ctx, ctx.gather_device, torch, d_q, nq, k, r, filt, params, want_idx)
            assert (ref[1] == k).sum() > nq // 4                  # the estimate sizes are really reached
            first = _run(ctx, ctx.gather_device_noalloc, torch, d_q, nq, k, r, filt, params, want_idx)
            second = _run(ctx, ctx.gather_device_noalloc, torch, d_q, nq, k, r, filt, params, want_idx)
            _same(ref, first)
            _same(ref, second)
        # beyond the stated maximum: a clear error, nothing launched
        too_many = gi_b200.NOALLOC_MAX_QUERIES + 1
        with pytest.raises(gi_b200.GiError, match="GI_NOALLOC_MAX_QUERIES"):
            ctx.gather_device_noalloc(gi_b200.GI_MAP_CAUSTIC, d_q.data_ptr(), too_many, 50, 0.05, 0, params, None, None, None, None)
        # an empty batch is a no-op
        ctx.gather_device_noalloc(gi_b200.GI_MAP_CAUSTIC, None, 0, 50, 0.05, 0, params, None, None, None, None)
    finally:
        ctx.close()
