"""CPU checks of the renderer's counter-based RNG (csrc/rng.cuh, compiled for the host in tests/host_sim)."""
import ctypes
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sim():
    L = ctypes.CDLL(os.path.join(ROOT, "tests", "host_sim", "libisect_sim.so"))
    L.sim_rng_doubles.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p]
    return L


def block(ctr, key):
    c = (ctypes.c_uint32 * 4)(*ctr); k = (ctypes.c_uint32 * 2)(*key); o = (ctypes.c_uint32 * 4)()
    _sim().sim_philox_block(c, k, o)
    return list(o)


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    assert block([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert block([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert block([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def doubles(seed, index, stage, sample, n):
    out = np.zeros(n)
    _sim().sim_rng_doubles(seed, index, stage, sample, n, out.ctypes.data)
    return out


def test_streams_are_distinct_and_uniform():
    """Every component of the key (seed, index, stage, sample) selects a different stream -- a regression test for
    a sampler whose per-sample streams coincided (means right, variance N times too high)."""
    base = doubles(7, 11, 2, 3, 64)
    assert np.array_equal(base, doubles(7, 11, 2, 3, 64))
    for other in (doubles(8, 11, 2, 3, 64), doubles(7, 12, 2, 3, 64), doubles(7, 11, 3, 3, 64), doubles(7, 11, 2, 4, 64),
                  doubles(7, 11 + (1 << 32), 2, 3, 64)):
        assert not np.any(other == base)
    # first draws of 4096 consecutive sample streams: uniform and uncorrelated with their neighbours
    first = np.array([doubles(1, 5, 2, s, 2) for s in range(4096)])
    assert np.all((first >= 0) & (first < 1))
    assert abs(first.mean() - 0.5) < 0.02 and abs(first.var() - 1 / 12) < 0.01
    assert abs(np.corrcoef(first[:-1, 0], first[1:, 0])[0, 1]) < 0.06
    assert abs(np.corrcoef(first[:, 0], first[:, 1])[0, 1]) < 0.06
