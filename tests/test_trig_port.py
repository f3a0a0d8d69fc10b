"""cos / sin of the descriptor angle (Features/ORBextractor.cpp:173-174 -> glibc cosf / sinf, float overloads).

The oracle restates glibc's algorithm (oracle/orb_oracle.cpp:glibc_sincosf_port); here it is pinned against the libm of the
machine, bit for bit, on a sample of the angle range (tools/probe/glibc_sincosf_check.c is the exhaustive form: every float of
[0, 2 pi], 0 mismatches).  The CUDA form of the same restatement is checked against it in the GPU test below.
"""
import numpy as np
import pytest

from oracle import orb_oracle as orc


def _sample():
    # every 61st float of [0, 2 pi] from 2^-13 up (17 M / 61 per binade), every float next to a quadrant boundary, all angles
    # k * factorPI that fastAtan2 can produce at 1/16 degree, and the range ends
    lo, hi = np.float32(2.0 ** -13).view(np.uint32), np.float32(6.2831855).view(np.uint32)
    a = np.arange(int(lo), int(hi) + 64, 61, dtype=np.uint32).view(np.float32)
    q = np.concatenate([np.float32(k * np.pi / 4).view(np.uint32).astype(np.int64) + np.arange(-2000, 2001) for k in range(1, 9)])
    q = q.astype(np.uint32).view(np.float32)
    d = (np.arange(0, 360 * 16 + 1, dtype=np.float32) / np.float32(16)) * np.float32(3.1415926535897932384626433832795 / 180.0)
    return np.concatenate([a, q, d, np.array([0.0, 1e-30, 2.0 ** -12, 0.78539816, 119.0], np.float32)])


def test_port_equals_libm():
    y = _sample()
    s0, c0 = orc.sincosf(y, 0)
    s1, c1 = orc.sincosf(y, 1)
    assert np.array_equal(s0.view(np.uint32), s1.view(np.uint32))
    assert np.array_equal(c0.view(np.uint32), c1.view(np.uint32))


@pytest.mark.gpu
def test_cuda_port_equals_oracle():
    from orb_slam_tracking_b200 import frontend as fe
    y = _sample()
    y = y[y < 119.0]
    E = fe.ORBextractor(500, 1.2, 8, 20, 7, 640, 480, max_batch=1)
    sn, cs = E.sincosf_probe(y)
    s0, c0 = orc.sincosf(y, 0)
    assert np.array_equal(sn.view(np.uint32), s0.view(np.uint32))
    assert np.array_equal(cs.view(np.uint32), c0.view(np.uint32))
    s1, c1 = orc.sincosf(y, 1)   # and, on this box, the machine's libm
    assert np.array_equal(sn.view(np.uint32), s1.view(np.uint32))
    assert np.array_equal(cs.view(np.uint32), c1.view(np.uint32))
