"""ctypes binding of oracle/liborb_oracle.so.

TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs import this module.  See orb_oracle.cpp for what each entry restates.
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_LIB = None

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
     ("octave", "<i4"), ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28


def build(force=False):
    so = _DIR / "liborb_oracle.so"
    srcs = [_DIR / "orb_oracle.cpp", _DIR / "cv_prims.hpp"]
    if force or not so.exists() or any(s.exists() and s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        subprocess.check_call(["make", "-C", str(_DIR), "liborb_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = _DIR / "liborb_oracle.so"
        if not so.exists():
            build()
        L = C.CDLL(str(so))
        u8p, ip, fp = C.POINTER(C.c_ubyte), C.POINTER(C.c_int), C.POINTER(C.c_float)
        vp = C.c_void_p
        L.orc_extractor_create.restype = vp
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_destroy.argtypes = [vp]
        L.orc_extractor_set_trig_mode.argtypes = [vp, C.c_int]
        L.orc_extractor_set_policy.argtypes = [vp, C.c_int]
        L.orc_retain_best.argtypes = [vp, C.c_int, C.c_int, vp]
        L.orc_sincosf.argtypes = [C.c_int, vp, C.c_longlong, vp, vp]
        L.orc_extractor_tables.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orc_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp]
        L.orc_level_size.argtypes = [vp, C.c_int, vp, vp]
        L.orc_get_level.argtypes = [vp, C.c_int, C.c_int, vp]
        L.orc_get_blurred.argtypes = [vp, C.c_int, vp]
        L.orc_get_candidates.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orc_get_level_keypoints.argtypes = [vp, C.c_int, vp, vp, C.c_int]
        L.orc_distribute.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_resize_linear.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int]
        L.orc_border101.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int]
        L.orc_fast.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_blur7.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_undistort.argtypes = [vp, C.c_int, vp, vp, vp]
        L.orc_bgr2gray.argtypes = [vp, C.c_int, vp]
        L.orc_descriptor_distance.argtypes = [vp, vp]
        L.orc_image_bounds.argtypes = [C.c_int, C.c_int, vp, vp, vp]
        L.orc_frame_create.restype = vp
        L.orc_frame_create.argtypes = [vp, vp, C.c_int, vp, vp, vp]
        L.orc_frame_destroy.argtypes = [vp]
        L.orc_frame_keys_un.argtypes = [vp, vp]
        L.orc_frame_grid.argtypes = [vp, vp, vp]
        L.orc_features_in_area.argtypes = [vp, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, vp, C.c_int]
        L.orc_search_for_initialization.argtypes = [vp, vp, C.c_float, C.c_int, C.c_int, vp, vp]
        L.orc_run_stream.restype = C.c_longlong
        L.orc_run_stream.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_int, vp, vp, C.c_int, C.c_float, C.c_int, vp]
        L.orc_check_homography.argtypes = [vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_float, vp, vp]
        L.orc_check_fundamental.argtypes = [vp, vp, vp, C.c_int, vp, C.c_int, C.c_float, vp, vp]
        L.orc_ransac_best.argtypes = [vp, C.c_int]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


# ------------------------------------------------------------------ primitives
def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border101(src, b):
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    dst = np.empty((h + 2 * b, w + 2 * b), np.uint8)
    lib().orc_border101(_p(src), w, h, src.strides[0], _p(dst), w + 2 * b, b)
    return dst


def fast(img, th, cap=1 << 20):
    img = np.ascontiguousarray(img, np.uint8)
    xs, ys, sc = (np.empty(cap, np.int32) for _ in range(3))
    n = lib().orc_fast(_p(img), img.shape[1], img.shape[0], img.strides[0], th, _p(xs), _p(ys), _p(sc), cap)
    return xs[:n].copy(), ys[:n].copy(), sc[:n].copy()


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(float(y), float(x)))


def blur7(src):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty_like(src)
    lib().orc_blur7(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dst.strides[0])
    return dst


def undistort(pts, K4, D4):
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    K4 = np.ascontiguousarray(K4, np.float32)
    D4 = np.ascontiguousarray(D4, np.float32)
    out = np.empty_like(pts)
    lib().orc_undistort(_p(pts), len(pts), _p(K4), _p(D4), _p(out))
    return out


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty(bgr.shape[:2], np.uint8)
    lib().orc_bgr2gray(_p(bgr), out.size, _p(out))
    return out


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return int(lib().orc_descriptor_distance(_p(a), _p(b)))


def distribute(xyr, min_x, max_x, min_y, max_y, n_features):
    xyr = np.ascontiguousarray(xyr, np.float32).reshape(-1, 3)
    out = np.empty((max(len(xyr), 1), 3), np.float32)
    n = lib().orc_distribute(_p(xyr), len(xyr), min_x, max_x, min_y, max_y, n_features, _p(out), len(out))
    if n < 0:
        raise RuntimeError("oracle distribute status %d" % n)
    return out[:n].copy()


def image_bounds(W, H, K4, D4):
    K4 = np.ascontiguousarray(K4, np.float32)
    D4 = np.ascontiguousarray(D4, np.float32)
    b = np.zeros(4, np.int32)
    lib().orc_image_bounds(W, H, _p(K4), _p(D4), _p(b))
    return b


# ------------------------------------------------------------------ extractor
class Extractor:
    """Mirror of ORB_SLAM_Tracking::ORBextractor (reference Features/ORBextractor.hpp:55-158)."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th, min_th, trig_mode=2, policy=0):
        """policy 0: ComputeKeyPointsOctTree (the reference's operator()); 1: ComputeKeyPointsOld (its commented-out alternative)."""
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = lib().orc_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        if not self._h:
            raise ValueError("scaleFactor == 1.0 but nlevels is not 1")
        lib().orc_extractor_set_trig_mode(self._h, trig_mode)
        lib().orc_extractor_set_policy(self._h, policy)
        f = lambda: np.zeros(nlevels, np.float32)
        self.scale, self.inv_scale, self.sigma2, self.inv_sigma2 = f(), f(), f(), f()
        self.per_level = np.zeros(nlevels, np.int32)
        self.umax = np.zeros(16, np.int32)
        lib().orc_extractor_tables(self._h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2),
                                   _p(self.inv_sigma2), _p(self.per_level), _p(self.umax))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def __call__(self, img, lapping=(0, 0)):
        """-> (keypoints[KP_DTYPE], descriptors[n,32] u8, monoIndex)"""
        assert img.dtype == np.uint8 and img.ndim == 2
        H, W = img.shape
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        lap = np.asarray(lapping, np.int32)
        mono = C.c_int(0)
        n = lib().orc_extract(self._h, _p(img), W, H, img.strides[0], _p(lap), _p(kps), _p(desc), cap,
                              C.addressof(mono))
        if n < 0:
            raise RuntimeError("oracle extract status %d" % n)
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy(), mono.value

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self._h, level, C.addressof(w), C.addressof(h))
        return w.value, h.value

    def level(self, level, padded=False):
        w, h = self.level_size(level)
        out = np.empty((h + 38, w + 38) if padded else (h, w), np.uint8)
        lib().orc_get_level(self._h, level, int(padded), _p(out))
        return out

    def blurred(self, level):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        return out if lib().orc_get_blurred(self._h, level, _p(out)) else None

    def candidates(self, level):
        n = lib().orc_get_candidates(self._h, level, None, 0)
        out = np.empty((max(n, 1), 3), np.float32)
        lib().orc_get_candidates(self._h, level, _p(out), n)
        return out[:n]

    def level_keypoints(self, level):
        n = lib().orc_get_level_keypoints(self._h, level, None, None, 0)
        kps = np.zeros(max(n, 1), KP_DTYPE)
        desc = np.zeros((max(n, 1), 32), np.uint8)
        lib().orc_get_level_keypoints(self._h, level, _p(kps), _p(desc), n)
        return kps[:n], desc[:n]


# ------------------------------------------------------------------ frame + matcher
class Frame:
    """Mirror of the Frame constructor glue (reference SlamTypes/Frame.cpp:40-206)."""

    def __init__(self, kps, desc, K4, D4, bounds):
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.N = len(self.kps)
        K4 = np.ascontiguousarray(K4, np.float32)
        D4 = np.ascontiguousarray(D4, np.float32)
        b = np.ascontiguousarray(bounds, np.int32)
        self._h = lib().orc_frame_create(_p(self.kps), _p(self.desc), self.N, _p(K4), _p(D4), _p(b))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_frame_destroy(self._h)
            self._h = None

    def keys_un(self):
        out = np.zeros(max(self.N, 1), KP_DTYPE)
        n = lib().orc_frame_keys_un(self._h, _p(out))
        return out[:n]

    def grid(self):
        off = np.zeros(64 * 48 + 1, np.uint32)
        idx = np.zeros(max(self.N, 1), np.uint32)
        n = lib().orc_frame_grid(self._h, _p(off), _p(idx))
        return off, idx[:n]

    def features_in_area(self, x, y, r, min_level=-1, max_level=-1):
        out = np.zeros(max(self.N, 1), np.int32)
        n = lib().orc_features_in_area(self._h, x, y, r, min_level, max_level, _p(out), len(out))
        return out[:n].copy()


def search_for_initialization(F1, F2, nnratio=0.9, check_ori=True, window=100):
    m12 = np.full(max(F1.N, 1), -1, np.int32)
    cnt = np.zeros(3, np.int32)
    n = lib().orc_search_for_initialization(F1._h, F2._h, nnratio, int(check_ori), window, _p(m12), _p(cnt))
    return n, m12[:F1.N].copy(), cnt


def run_stream(imgs, nfeatures, scale_factor, nlevels, ini_th, min_th, K4, D4, window=100, nnratio=0.9,
               threads=1):
    """CPU baseline: extract every frame, glue, match pairs (2k, 2k+1). -> (matches, keypoints)"""
    imgs = np.ascontiguousarray(imgs, np.uint8)
    n, H, W = imgs.shape
    K4 = np.ascontiguousarray(K4, np.float32)
    D4 = np.ascontiguousarray(D4, np.float32)
    tot = C.c_longlong(0)
    m = lib().orc_run_stream(_p(imgs), n, W, H, nfeatures, scale_factor, nlevels, ini_th, min_th, _p(K4), _p(D4),
                             window, nnratio, threads, C.addressof(tot))
    return int(m), int(tot.value)


# ------------------------------------------------------------------ Initializer scoring (row N2)
def _hyp(M):
    M = np.ascontiguousarray(M, np.float32).reshape(-1, 9)
    return M


def check_homography(kps1_un, kps2_un, matches12, H21, H12, sigma=1.0):
    """Initializer::CheckHomography for each hypothesis -> (scores[nHyp], inliers[nHyp, N])."""
    k1, k2 = np.ascontiguousarray(kps1_un, KP_DTYPE), np.ascontiguousarray(kps2_un, KP_DTYPE)
    m = np.ascontiguousarray(matches12, np.int32)
    A, B = _hyp(H21), _hyp(H12)
    N = int((m >= 0).sum())
    scores = np.zeros(len(A), np.float32)
    inl = np.zeros((len(A), N), np.uint8)
    lib().orc_check_homography(_p(k1), _p(k2), _p(m), len(m), _p(A), _p(B), len(A), sigma, _p(scores), _p(inl))
    return scores, inl


def check_fundamental(kps1_un, kps2_un, matches12, F21, sigma=1.0):
    k1, k2 = np.ascontiguousarray(kps1_un, KP_DTYPE), np.ascontiguousarray(kps2_un, KP_DTYPE)
    m = np.ascontiguousarray(matches12, np.int32)
    A = _hyp(F21)
    N = int((m >= 0).sum())
    scores = np.zeros(len(A), np.float32)
    inl = np.zeros((len(A), N), np.uint8)
    lib().orc_check_fundamental(_p(k1), _p(k2), _p(m), len(m), _p(A), len(A), sigma, _p(scores), _p(inl))
    return scores, inl


def ransac_best(scores):
    s = np.ascontiguousarray(scores, np.float32)
    return int(lib().orc_ransac_best(_p(s), len(s)))


def sincosf(y, which=0):
    """sin / cos of float32 values by the restated glibc sinf / cosf (which=0) or by the machine's libm (which=1)."""
    y = np.ascontiguousarray(y, np.float32)
    sn, cs = np.empty_like(y), np.empty_like(y)
    lib().orc_sincosf(which, y.ctypes.data, y.size, sn.ctypes.data, cs.ctypes.data)
    return sn, cs


def retain_best(response, n_points):
    """Indices kept by cv::KeyPointsFilter::retainBest (restated, cv_prims.hpp) in their final order."""
    r = np.ascontiguousarray(response, np.float32)
    idx = np.empty(len(r), np.int32)
    n = lib().orc_retain_best(r.ctypes.data, len(r), int(n_points), idx.ctypes.data)
    return idx[:n].copy()
