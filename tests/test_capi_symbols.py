"""CPU-only: liborbx.so builds for sm_100a, loads without a GPU and exports every symbol that
include/orbx.h declares; the ctypes table covers the same set.  No compute call is made here, except the
host-only entry points (constructor tables, 32-byte Hamming distance)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "orbx.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from orb_slam_tracking_b200 import build, capi
    build.build()
    L = C.CDLL(str(capi.LIB_PATH))
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), n
    assert sorted(capi.SIGNATURES) == names
    capi.lib()


def test_host_only_entry_points(oracle):
    from orb_slam_tracking_b200 import capi
    L = capi.lib()
    for nf, sf, nl in ((1000, 1.2, 8), (2000, 1.2, 8), (1250, 1.2, 8), (500, 1.5, 4), (8000, 1.2, 8)):
        f = lambda: np.zeros(nl, np.float32)
        s, i, s2, i2 = f(), f(), f(), f()
        npl, um = np.zeros(nl, np.int32), np.zeros(16, np.int32)
        assert L.orbx_compute_level_tables(nf, sf, nl, capi.ptr(s), capi.ptr(i), capi.ptr(s2), capi.ptr(i2),
                                           capi.ptr(npl), capi.ptr(um)) == nl
        R = oracle.Extractor(nf, sf, nl, 20, 7)
        for a, b in ((s, R.scale), (i, R.inv_scale), (s2, R.sigma2), (i2, R.inv_sigma2)):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert np.array_equal(npl, R.per_level) and np.array_equal(um, R.umax)
    assert L.orbx_compute_level_tables(1000, 1.0, 8, None, None, None, None, None, None) == capi.ERR_CTOR
    rng = np.random.default_rng(0)
    for _ in range(100):
        a = rng.integers(0, 256, 32, dtype=np.uint8)
        b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert L.orbx_descriptor_distance(capi.ptr(a), capi.ptr(b)) == oracle.descriptor_distance(a, b)


def test_create_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return
    from orb_slam_tracking_b200 import capi, frontend
    try:
        frontend.ORBextractor(1000, 1.2, 8, 20, 7, 640, 480)
    except capi.OrbxError as e:
        assert e.code == capi.ERR_CUDA
    else:
        raise AssertionError("orbx_create succeeded without a CUDA device: a CPU fallback must not exist")
