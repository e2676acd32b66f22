"""CPU suite: the oracle restatement of the photon gather (oracle/restate_gather.cpp) against the reference's own
R3Kdtree::FindClosestQuick + EstimateRadiance on reference-traced cornell maps (tests/golden/cornell.gather.npz).
Bar: identical kNN index sets and counts; radiance within 1e-12 relative (summation order only)."""
import os

import numpy as np
import pytest

from gi_testlib import ROOT, read_photons, synthetic_photons, GI_QUERY

GOLDEN = os.path.join(ROOT, "tests", "golden")
CASES = [("caustic", 225), ("global", 50), ("caustic", 40)]


@pytest.fixture(scope="module")
def fixture():
    return np.load(os.path.join(GOLDEN, "cornell.gather.npz"))


@pytest.mark.parametrize("mp,k", CASES)
@pytest.mark.parametrize("filt", [0, 1, 2])
def test_oracle_gather_matches_reference(oracle, fixture, mp, k, filt):
    ph = read_photons(os.path.join(GOLDEN, f"cornell.{mp}.gipm"))
    key = f"{mp}_{k}_{filt}"
    r = float(fixture[key + "_r"][0])
    idx, cnt, rgb = oracle.gather(ph, fixture["queries"], k, r, filt, threads=4)
    assert np.array_equal(cnt, fixture[key + "_count"])
    assert np.array_equal(idx, fixture[f"{mp}_{k}_idx"])
    ref = fixture[key + "_rgb"]
    assert np.allclose(rgb, ref, rtol=1e-12, atol=1e-300)
    assert (ref.sum(1) > 0).sum() > 100        # the fixture actually exercises the estimate


def test_oracle_gather_edge_cases(oracle):
    q = np.zeros(3, dtype=GI_QUERY)
    q["n"] = (0, 0, 1); q["bounce"] = (0, 0, 1); q["cos_theta"] = 1.0; q["kd"] = 1.0
    q["p"] = [(0, 0, 1), (5, 5, 5), (0.999, 0.999, 1)]
    # empty map
    idx, cnt, rgb = oracle.gather(synthetic_photons(0, 1), q, 10, 0.5)
    assert (cnt == 0).all() and (idx == -1).all() and (rgb == 0).all()
    # fewer photons in range than k; a far query finds nothing
    ph = synthetic_photons(500, 2)
    idx, cnt, rgb = oracle.gather(ph, q, 50, 0.2)
    assert cnt[1] == 0 and (idx[1] == -1).all()
    assert 0 < cnt[0] < 50 and (np.diff(idx[0][:cnt[0]]) > 0).all()
    # exact ties: duplicated photons -> lower index wins
    dup = np.concatenate([ph[:100], ph[:100]])
    idx, cnt, _ = oracle.gather(dup, q, 1, 10.0)
    assert (idx[:, 0] < 100).all()
