"""ctypes binding of oracle/_ref/libref_orb.so: the reference's OWN sources (ORBextractor.cpp,
ORBmatcher.cpp, Frame.cpp, DBoW2 FORB.cpp) compiled in place by oracle/build_ref.sh against the
header stand-in oracle/refshim.

TEST INFRASTRUCTURE.  Only tests/, tools/gen_golden.py and bench.py's cpu_baseline / --impl
reference legs may import this module.
"""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from .orb_oracle import KP_DTYPE

_DIR = Path(__file__).resolve().parent
SO = _DIR / "_ref" / "libref_orb.so"
_LIB = None


def build():
    """Needs /root/reference; a no-op (and `available()` stays False) without it."""
    srcs = [_DIR / "ref_driver.cpp", _DIR / "ref_init_driver.cpp", _DIR / "ref_old_driver.cpp", _DIR / "build_ref.sh", _DIR / "cv_prims.hpp", _DIR / "refshim" / "opencv2" / "opencv.hpp"]
    if SO.exists() and all(s.stat().st_mtime <= SO.stat().st_mtime for s in srcs):
        return SO
    subprocess.run(["bash", str(_DIR / "build_ref.sh")], check=True, stdout=subprocess.DEVNULL)
    return SO


def available():
    return SO.exists()


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(str(SO))
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.ref_extractor_create.restype = vp
        L.ref_extractor_create.argtypes = [ci, cf, ci, ci, ci]
        L.ref_extractor_destroy.argtypes = [vp]
        L.ref_extractor_tables.argtypes = [vp] * 6
        L.ref_extract.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp, ci, vp]
        L.ref_get_level.argtypes = [vp, ci, ci, vp, vp, vp]
        L.ref_frame_create.restype = vp
        L.ref_frame_create.argtypes = [vp, vp, ci, ci, ci, vp, vp, ci]
        L.ref_frame_destroy.argtypes = [vp]
        L.ref_frame_n.argtypes = [vp]
        L.ref_frame_bounds.argtypes = [vp]
        L.ref_frame_get.argtypes = [vp, vp, vp, vp]
        L.ref_frame_grid.argtypes = [vp, vp, vp]
        L.ref_features_in_area.argtypes = [vp, cf, cf, cf, ci, ci, vp, ci]
        L.ref_search_for_initialization.argtypes = [vp, vp, cf, ci, ci, vp]
        L.ref_forb_distance.restype = C.c_double
        L.ref_forb_distance.argtypes = [vp, vp]
        L.ref_run_stream.restype = C.c_longlong
        L.ref_run_stream.argtypes = [vp, ci, ci, ci, ci, cf, ci, ci, ci, vp, vp, ci, cf, ci, vp]
        if hasattr(L, "ref_old_extract"):
            L.ref_old_extractor_create.restype = vp
            L.ref_old_extractor_create.argtypes = [ci, cf, ci, ci, ci]
            L.ref_old_extractor_destroy.argtypes = [vp]
            L.ref_old_extract.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp, ci, vp]
        if hasattr(L, "ref_check_homography"):
            L.ref_check_homography.argtypes = [vp, ci, vp, ci, vp, vp, vp, ci, cf, vp, vp]
            L.ref_check_fundamental.argtypes = [vp, ci, vp, ci, vp, vp, ci, cf, vp, vp]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Extractor:
    """ORB_SLAM_Tracking::ORBextractor (Features/ORBextractor.hpp:55-158) of the reference."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().ref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_extractor_destroy(self._h)
            self._h = None

    def tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        q = np.zeros(n, np.int32)
        lib().ref_extractor_tables(self._h, *[_p(a) for a in f], _p(q))
        return (*f, q)

    def __call__(self, img, lapping=(0, 0)):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = max(4 * self.nfeatures, 64)
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = C.c_int(0)
        n = lib().ref_extract(self._h, _p(img), w, h, img.strides[0], lapping[0], lapping[1], _p(kps), _p(desc), cap, C.byref(mono))
        assert n >= -1, "capacity"
        n = max(n, 0)
        return kps[:n].copy(), desc[:n].copy(), mono.value

    def level(self, level, padded=False):
        w, h = C.c_int(0), C.c_int(0)
        lib().ref_get_level(self._h, level, int(padded), None, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        lib().ref_get_level(self._h, level, int(padded), _p(out), C.byref(w), C.byref(h))
        return out


class OldPolicyExtractor:
    """The reference's ORBextractor with operator() calling ComputeKeyPointsOld (Features/ORBextractor.cpp:1207-1496) instead
    of ComputeKeyPointsOctTree: the reference's own code, one call swapped at build time (oracle/build_ref.sh)."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th):
        self.nfeatures = nfeatures
        self._h = lib().ref_old_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_old_extractor_destroy(self._h)
            self._h = None

    def __call__(self, img, lapping=(0, 0)):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = max(4 * self.nfeatures, 64)
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mono = C.c_int(0)
        n = lib().ref_old_extract(self._h, _p(img), w, h, img.strides[0], lapping[0], lapping[1], _p(kps), _p(desc), cap, C.byref(mono))
        assert n >= -1, "capacity"
        n = max(n, 0)
        return kps[:n].copy(), desc[:n].copy(), mono.value


class Frame:
    """ORB_SLAM_Tracking::Frame (SlamTypes/Frame.hpp:20-136) built by its real constructor."""

    def __init__(self, extractor, img, K4, D4, first=True):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        K4 = np.ascontiguousarray(K4, np.float32)
        D4 = np.ascontiguousarray(D4, np.float32)
        self._e = extractor
        self._h = lib().ref_frame_create(extractor._h, _p(img), w, h, img.strides[0], _p(K4), _p(D4), int(first))
        self.n = lib().ref_frame_n(self._h)
        b = np.zeros(4, np.int32)
        lib().ref_frame_bounds(_p(b))
        self.bounds = b

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_frame_destroy(self._h)
            self._h = None

    def get(self):
        k, ku, d = np.zeros(self.n, KP_DTYPE), np.zeros(self.n, KP_DTYPE), np.zeros((self.n, 32), np.uint8)
        lib().ref_frame_get(self._h, _p(k), _p(ku), _p(d))
        return k, ku, d

    def grid(self):
        off, idx = np.zeros(64 * 48 + 1, np.uint32), np.zeros(max(self.n, 1), np.uint32)
        lib().ref_frame_grid(self._h, _p(off), _p(idx))
        return off, idx[:off[-1]]

    def features_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.n, 1), np.int32)
        n = lib().ref_features_in_area(self._h, x, y, r, min_level, max_level, _p(out), len(out))
        return out[:n].copy()


def search_for_initialization(F1, F2, nnratio=0.9, check_ori=True, window=100):
    m = np.full(max(F1.n, 1), -1, np.int32)
    n = lib().ref_search_for_initialization(F1._h, F2._h, nnratio, int(check_ori), window, _p(m))
    return n, m[:F1.n].copy()


def forb_distance(a, b):
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib().ref_forb_distance(_p(a), _p(b))


def run_stream(imgs, nfeatures, scale_factor, nlevels, ini_th, min_th, K4, D4, window=100, nnratio=0.9, threads=1):
    """CPU baseline on the reference's own code: Frame x 2 + SearchForInitialization per pair. -> (matches, keypoints)"""
    imgs = np.ascontiguousarray(imgs, np.uint8)
    n, H, W = imgs.shape
    K4 = np.ascontiguousarray(K4, np.float32)
    D4 = np.ascontiguousarray(D4, np.float32)
    tot = C.c_longlong(0)
    m = lib().ref_run_stream(_p(imgs), n, W, H, nfeatures, scale_factor, nlevels, ini_th, min_th, _p(K4), _p(D4), window, nnratio,
                             threads, C.byref(tot))
    return int(m), int(tot.value)


def check_homography(kps1_un, kps2_un, matches12, H21, H12, sigma=1.0):
    """The reference's own Initializer::CheckHomography (Initializer.cpp:268-353) per hypothesis."""
    k1, k2 = np.ascontiguousarray(kps1_un, KP_DTYPE), np.ascontiguousarray(kps2_un, KP_DTYPE)
    m = np.ascontiguousarray(matches12, np.int32)
    A = np.ascontiguousarray(H21, np.float32).reshape(-1, 9)
    B = np.ascontiguousarray(H12, np.float32).reshape(-1, 9)
    N = int((m >= 0).sum())
    scores = np.zeros(len(A), np.float32)
    inl = np.zeros((len(A), N), np.uint8)
    lib().ref_check_homography(_p(k1), len(k1), _p(k2), len(k2), _p(m), _p(A), _p(B), len(A), sigma, _p(scores), _p(inl))
    return scores, inl


def check_fundamental(kps1_un, kps2_un, matches12, F21, sigma=1.0):
    k1, k2 = np.ascontiguousarray(kps1_un, KP_DTYPE), np.ascontiguousarray(kps2_un, KP_DTYPE)
    m = np.ascontiguousarray(matches12, np.int32)
    A = np.ascontiguousarray(F21, np.float32).reshape(-1, 9)
    N = int((m >= 0).sum())
    scores = np.zeros(len(A), np.float32)
    inl = np.zeros((len(A), N), np.uint8)
    lib().ref_check_fundamental(_p(k1), len(k1), _p(k2), len(k2), _p(m), _p(A), len(A), sigma, _p(scores), _p(inl))
    return scores, inl
