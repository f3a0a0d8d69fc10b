"""Python mirror of the reference's front-end classes over the C ABI (used by tests and bench.py;
the C++ drop-in headers live in include/).  Names follow the reference:
  ORBextractor   Features/ORBextractor.hpp:55-158
  Frame          SlamTypes/Frame.hpp:20-136 (constructor glue, GetFeaturesInArea)
  ORBmatcher     Features/ORBmatcher.hpp:7-63
  Stream         device-resident extract -> glue -> SearchForInitialization over a batch of frames
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import GRID_CELLS, KP_DTYPE, check, lib, ptr


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) bound to one image shape."""

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, width, height, max_batch=1, device=0):
        st = C.c_int(0)
        self._h = lib().orbx_create(device, width, height, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST,
                                    max_batch, C.addressof(st))
        if not self._h:
            raise capi.OrbxError(st.value, "orbx_create")
        self.nfeatures, self.nlevels, self.width, self.height, self.max_batch = nfeatures, nlevels, width, height, max_batch
        self.scaleFactor = scaleFactor
        f = lambda: np.zeros(nlevels, np.float32)
        self._scale, self._inv, self._s2, self._is2 = f(), f(), f(), f()
        self._npl = np.zeros(nlevels, np.int32)
        self.umax = np.zeros(16, np.int32)
        check(lib().orbx_get_level_tables(self._h, ptr(self._scale), ptr(self._inv), ptr(self._s2), ptr(self._is2),
                                          ptr(self._npl), ptr(self.umax)))
        self.cap = max(nfeatures, int(self._npl.sum()))

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_destroy(self._h)
            self._h = None

    __del__ = close

    # accessors, Features/ORBextractor.hpp:87-108 (copies, like the reference)
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return self.scaleFactor
    def GetScaleFactors(self): return self._scale.copy()
    def GetInverseScaleFactors(self): return self._inv.copy()
    def GetScaleSigmaSquares(self): return self._s2.copy()
    def GetInverseScaleSigmaSquares(self): return self._is2.copy()
    def GetNumFeaturesPerLevel(self): return self._npl.copy()

    def set_trig_mode(self, mode):
        check(lib().orbx_set_trig_mode(self._h, mode), self._h)

    def sincosf_probe(self, y):
        """sin / cos of float32 angles as the descriptor kernel computes them in trig mode 2 (orbx_sincosf_probe)."""
        y = np.ascontiguousarray(y, np.float32)
        sn, cs = np.empty_like(y), np.empty_like(y)
        check(lib().orbx_sincosf_probe(self._h, ptr(y), y.size, ptr(sn), ptr(cs)), self._h)
        return sn, cs

    def set_keypoint_policy(self, policy):
        """0: ComputeKeyPointsOctTree (default), 1: ComputeKeyPointsOld (Features/ORBextractor.cpp:1207-1496)."""
        check(lib().orbx_set_keypoint_policy(self._h, int(policy)), self._h)

    def set_option(self, name, value):
        """Schedule options (orbx_set_option): they never change results."""
        check(lib().orbx_set_option(self._h, name.encode(), int(value)), self._h)

    def hamming_best2(self, a, b, repeats=1):
        """Raw all-pairs Hamming: -> (best_idx[nA], best_dist[nA], second_dist[nA], kernel_ms)."""
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        bi, bd, sd = (np.zeros(len(a), np.int32) for _ in range(3))
        ms = C.c_float(0)
        check(lib().orbx_hamming_best2(self._h, ptr(a), len(a), ptr(b), len(b), ptr(bi), ptr(bd), ptr(sd), repeats,
                                       C.addressof(ms)), self._h)
        return bi, bd, sd, ms.value

    def __call__(self, image, mask=None, vLappingArea=(0, 0)):
        """-> (keypoints[KP_DTYPE], descriptors[n,32], monoIndex).  `mask` is ignored as in the
        reference (Features/ORBextractor.cpp:1537)."""
        if image is None or image.size == 0:
            return -1
        assert image.dtype == np.uint8 and image.ndim == 2 and image.shape == (self.height, self.width)
        assert image.strides[1] == 1
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        lap = np.asarray(vLappingArea, np.int32)
        n, mono = C.c_int(0), C.c_int(0)
        check(lib().orbx_extract(self._h, ptr(image), image.strides[0], ptr(lap), ptr(kps), ptr(desc), self.cap,
                                 C.addressof(n), C.addressof(mono)), self._h)
        return kps[:n.value].copy(), desc[:n.value].copy(), mono.value

    def extract_batch(self, images, vLappingArea=(0, 0)):
        n_img = len(images)
        images = [np.ascontiguousarray(im, np.uint8) for im in images]
        arr = (C.c_void_p * n_img)(*[im.ctypes.data for im in images])
        kps = np.zeros((n_img, self.cap), KP_DTYPE)
        desc = np.zeros((n_img, self.cap, 32), np.uint8)
        lap = np.asarray(vLappingArea, np.int32)
        n = np.zeros(n_img, np.int32)
        mono = np.zeros(n_img, np.int32)
        check(lib().orbx_extract_batch(self._h, C.cast(arr, C.c_void_p), self.width, n_img, ptr(lap), ptr(kps),
                                       ptr(desc), self.cap, ptr(n), ptr(mono)), self._h)
        return [(kps[i, :n[i]].copy(), desc[i, :n[i]].copy(), int(mono[i])) for i in range(n_img)]

    # inspection of the last batch
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        check(lib().orbx_get_level_size(self._h, level, C.addressof(w), C.addressof(h)))
        return w.value, h.value

    def pyramid_level(self, level, padded=False, slot=0):
        """mvImagePyramid[level] (Features/ORBextractor.hpp:111); padded adds the 19-px reflect frame."""
        w, h = self.level_size(level)
        b = 19 if padded else 0
        out = np.empty((h + 2 * b, w + 2 * b), np.uint8)
        check(lib().orbx_get_pyramid_level(self._h, slot, level, int(padded), ptr(out), out.strides[0]), self._h)
        return out

    def blurred_level(self, level, slot=0):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        check(lib().orbx_get_blurred_level(self._h, slot, level, ptr(out), out.strides[0]), self._h)
        return out

    def candidates(self, level, slot=0):
        n = check(lib().orbx_get_candidates(self._h, slot, level, None, 0), self._h)
        out = np.zeros((max(n, 1), 3), np.float32)
        check(lib().orbx_get_candidates(self._h, slot, level, ptr(out), n), self._h)
        return out[:n]

    def level_keypoints(self, level, slot=0):
        cap = int(self._npl[level]) + 1
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = check(lib().orbx_get_level_keypoints(self._h, slot, level, ptr(kps), ptr(desc), cap), self._h)
        return kps[:n].copy(), desc[:n].copy()

    def set_camera(self, K4, D4, bounds=None):
        K4 = np.ascontiguousarray(K4, np.float32)
        D4 = np.ascontiguousarray(D4, np.float32)
        b = None if bounds is None else np.ascontiguousarray(bounds, np.int32)
        out = np.zeros(4, np.int32)
        check(lib().orbx_set_camera(self._h, ptr(K4), ptr(D4), ptr(b), ptr(out)), self._h)
        return out


class Frame:
    """Frame(im, extractor, K4, D4): extract -> undistort -> grid, as the reference constructor
    (SlamTypes/Frame.cpp:40-82).  The extractor's camera must have been set (statics in the reference)."""

    def __init__(self, extractor, im=None, kps=None, desc=None, fused=False, vLappingArea=(0, 0)):
        self.ex = extractor
        if im is not None and fused:
            # the single-call form (orbx_frame_construct): one CUDA graph, one synchronisation
            cap = extractor.cap
            k, d = np.zeros(cap, KP_DTYPE), np.zeros((cap, 32), np.uint8)
            ku, off, idx = np.zeros(cap, KP_DTYPE), np.zeros(GRID_CELLS + 1, np.uint32), np.zeros(cap, np.uint32)
            n, mono = C.c_int(0), C.c_int(0)
            lap = np.asarray(vLappingArea, np.int32)
            check(lib().orbx_frame_construct(extractor._h, ptr(im), im.strides[0], ptr(lap), ptr(k), ptr(d), cap, C.addressof(n),
                                             C.addressof(mono), ptr(ku), ptr(off), ptr(idx)), extractor._h)
            self.N, self.monoIndex = n.value, mono.value
            self.mvKeys, self.mDescriptors, self.mvKeysUn = k[:self.N].copy(), d[:self.N].copy(), ku[:self.N].copy()
            self.grid_offsets, self.grid_indices = off, idx[:int(off[-1])].copy()
            return
        if im is not None:
            kps, desc, _ = extractor(im, vLappingArea=vLappingArea)
        self.mvKeys = np.ascontiguousarray(kps, KP_DTYPE)
        self.mDescriptors = np.ascontiguousarray(desc, np.uint8)
        self.N = len(self.mvKeys)
        self.mvKeysUn = np.zeros(max(self.N, 1), KP_DTYPE)
        self.grid_offsets = np.zeros(GRID_CELLS + 1, np.uint32)
        self.grid_indices = np.zeros(max(self.N, 1), np.uint32)
        check(lib().orbx_frame_glue(extractor._h, ptr(self.mvKeys), self.N, ptr(self.mvKeysUn), ptr(self.grid_offsets),
                                    ptr(self.grid_indices)), extractor._h)
        self.mvKeysUn = self.mvKeysUn[:self.N]
        self.grid_indices = self.grid_indices[:int(self.grid_offsets[-1])]

    def _pod(self):
        f = capi.Frame()
        f.n = self.N
        f.keys_un = self.mvKeysUn.ctypes.data
        f.desc = self.mDescriptors.ctypes.data
        f.grid_offsets = self.grid_offsets.ctypes.data
        f.grid_indices = self.grid_indices.ctypes.data
        return f

    def mGrid(self, ix, iy):
        c = ix * capi.GRID_ROWS + iy
        return self.grid_indices[self.grid_offsets[c]:self.grid_offsets[c + 1]]

    def GetFeaturesInArea(self, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(self.N, 1), np.int32)
        f = self._pod()
        n = check(lib().orbx_features_in_area(self.ex._h, C.addressof(f), x, y, r, minLevel, maxLevel, ptr(out),
                                              len(out)), self.ex._h)
        return out[:n].copy()


class ORBmatcher:
    HISTO_LENGTH, TH_LOW, TH_HIGH = 30, 50, 100

    def __init__(self, nnratio=0.6, checkOri=True):
        self.mfNNratio, self.mbCheckOrientation = nnratio, checkOri

    def SearchForInitialization(self, F1, F2, windowSize=100):
        """-> (nmatches, vnMatches12, counters[3])"""
        m12 = np.full(max(F1.N, 1), -1, np.int32)
        nm = C.c_int(0)
        cnt = np.zeros(3, np.int32)
        f1, f2 = F1._pod(), F2._pod()
        check(lib().orbx_search_for_initialization(F1.ex._h, C.addressof(f1), C.addressof(f2), self.mfNNratio,
                                                   int(self.mbCheckOrientation), windowSize, ptr(m12),
                                                   C.addressof(nm), ptr(cnt)), F1.ex._h)
        return nm.value, m12[:F1.N].copy(), cnt

    def match_rows(self, F1, F2, windowSize, q0, q1):
        """Rows [q0, q1) of the pair's distance matrix as top lists -> (rows[n, 8] uint32, number of level-0 queries)."""
        f1, f2 = F1._pod(), F2._pod()
        rows = np.zeros((max(q1 - q0, 1), 8), np.uint32)
        nq = C.c_int(0)
        n = check(lib().orbx_match_rows(F1.ex._h, C.addressof(f1), C.addressof(f2), windowSize, q0, q1, ptr(rows),
                                        C.addressof(nq)), F1.ex._h)
        return rows[:n].copy(), nq.value

    def match_resolve(self, F1, F2, rows, windowSize=100):
        """The ordered resolve from a complete row table -> (nmatches, vnMatches12, counters[3])."""
        rows = np.ascontiguousarray(rows, np.uint32).reshape(-1, 8)
        m12 = np.full(max(F1.N, 1), -1, np.int32)
        nm = C.c_int(0)
        cnt = np.zeros(3, np.int32)
        f1, f2 = F1._pod(), F2._pod()
        check(lib().orbx_match_resolve(F1.ex._h, C.addressof(f1), C.addressof(f2), self.mfNNratio, int(self.mbCheckOrientation),
                                       windowSize, ptr(rows), len(rows), ptr(m12), C.addressof(nm), ptr(cnt)), F1.ex._h)
        return nm.value, m12[:F1.N].copy(), cnt

    @staticmethod
    def DescriptorDistance(a, b):
        a = np.ascontiguousarray(a, np.uint8)
        b = np.ascontiguousarray(b, np.uint8)
        return int(lib().orbx_descriptor_distance(ptr(a), ptr(b)))


class Initializer:
    """The scoring half of the reference's Initializer (Initialization/Initializer.hpp:57-85): CheckHomography and
    CheckFundamental over a set of hypotheses, and the selection rule of the RANSAC loops around them.  Estimating
    the hypotheses (cv::findHomography / findFundamentalMat in the reference) stays with the caller."""

    def __init__(self, referenceFrame, sigma=1.0, iterations=200):
        self.ex = referenceFrame.ex
        self.mvKeys1 = referenceFrame.mvKeysUn
        self.mSigma, self.mMaxIterations = sigma, iterations

    def _run(self, fn, currentFrame, matches12, mats):
        m12 = np.ascontiguousarray(matches12, np.int32)
        k2 = np.ascontiguousarray(currentFrame.mvKeysUn, KP_DTYPE)
        mats = [np.ascontiguousarray(M, np.float32).reshape(-1, 9) for M in mats]
        nh, N = len(mats[0]), int((m12 >= 0).sum())
        scores, inl = np.zeros(nh, np.float32), np.zeros((nh, N), np.uint8)
        nl, best = C.c_int(0), C.c_int(0)
        check(fn(self.ex._h, ptr(self.mvKeys1), len(self.mvKeys1), ptr(k2), len(k2), ptr(m12), *[ptr(M) for M in mats], nh,
                 self.mSigma, ptr(scores), ptr(inl), C.addressof(nl), C.addressof(best)), self.ex._h)
        assert nl.value == N
        return scores, inl, best.value

    def CheckHomography(self, currentFrame, matches12, H21, H12):
        """-> (scores[nHyp], vbMatchesInliers[nHyp, N], best)"""
        return self._run(lib().orbx_check_homography, currentFrame, matches12, [H21, H12])

    def CheckFundamental(self, currentFrame, matches12, F21):
        return self._run(lib().orbx_check_fundamental, currentFrame, matches12, [F21])


class Stream:
    """Batch of frames kept in HBM from image to match list (orbx_stream_*)."""

    def __init__(self, extractor, nnratio=0.9, checkOri=True, window=100):
        self.ex, self.nnratio, self.checkOri, self.window = extractor, nnratio, checkOri, window

    def upload(self, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8) if not isinstance(imgs, np.ndarray) else imgs
        assert imgs.flags.c_contiguous and imgs.shape[1:] == (self.ex.height, self.ex.width)
        check(lib().orbx_stream_upload(self.ex._h, ptr(imgs), imgs.shape[0]), self.ex._h)
        return imgs.shape[0]

    def upload_color(self, imgs, rgb=False):
        """[n, H, W, 3] interleaved BGR (or RGB) frames; Converter::toGray runs on the device."""
        imgs = np.ascontiguousarray(imgs, np.uint8)
        assert imgs.shape[1:] == (self.ex.height, self.ex.width, 3)
        check(lib().orbx_stream_upload_color(self.ex._h, ptr(imgs), imgs.shape[0], int(rgb)), self.ex._h)
        return imgs.shape[0]

    def run(self, n):
        check(lib().orbx_stream_run(self.ex._h, n, self.nnratio, int(self.checkOri), self.window), self.ex._h)

    def sync(self):
        check(lib().orbx_stream_sync(self.ex._h), self.ex._h)

    def download(self, n, full=True):
        cap = self.ex.cap
        out = {"n": np.zeros(n, np.int32), "nmatches": np.zeros(max(n // 2, 1), np.int32),
               "counters": np.zeros((max(n // 2, 1), 3), np.int32),
               "matches12": np.full((max(n // 2, 1), cap), -1, np.int32)}
        if full:
            out["kps"] = np.zeros((n, cap), KP_DTYPE)
            out["kps_un"] = np.zeros((n, cap), KP_DTYPE)
            out["desc"] = np.zeros((n, cap, 32), np.uint8)
        check(lib().orbx_stream_download(self.ex._h, n, ptr(out.get("kps")), ptr(out.get("kps_un")), ptr(out.get("desc")),
                                         cap, ptr(out["n"]), ptr(out["matches12"]), ptr(out["nmatches"]),
                                         ptr(out["counters"])), self.ex._h)
        return out

    def process(self, imgs, chunk=0, full=True):
        """Host frames in, everything out (orbx_stream_process: chunked, copies overlapped with the kernels)."""
        imgs = np.ascontiguousarray(imgs, np.uint8)
        assert imgs.shape[1:] == (self.ex.height, self.ex.width)
        n, cap = imgs.shape[0], self.ex.cap
        out = {"n": np.zeros(n, np.int32), "nmatches": np.zeros(max(n // 2, 1), np.int32),
               "counters": np.zeros((max(n // 2, 1), 3), np.int32),
               "matches12": np.full((max(n // 2, 1), cap), -1, np.int32)}
        if full:
            out["kps"] = np.zeros((n, cap), KP_DTYPE)
            out["kps_un"] = np.zeros((n, cap), KP_DTYPE)
            out["desc"] = np.zeros((n, cap, 32), np.uint8)
        check(lib().orbx_stream_process(self.ex._h, ptr(imgs), n, self.nnratio, int(self.checkOri), self.window, chunk,
                                        ptr(out.get("kps")), ptr(out.get("kps_un")), ptr(out.get("desc")), cap, ptr(out["n"]),
                                        ptr(out["matches12"]), ptr(out["nmatches"]), ptr(out["counters"])), self.ex._h)
        return out

    def process_async(self, imgs, chunk=0, full=True):
        """Streaming form (orbx_stream_process_async): returns a ticket; `wait(ticket)` returns that step's results.  Two
        steps may be in flight, so the upload of one overlaps the kernels and the download of the other."""
        imgs = np.ascontiguousarray(imgs, np.uint8)
        assert imgs.shape[1:] == (self.ex.height, self.ex.width)
        n, cap = imgs.shape[0], self.ex.cap
        out = {"n": np.zeros(n, np.int32), "nmatches": np.zeros(max(n // 2, 1), np.int32),
               "counters": np.zeros((max(n // 2, 1), 3), np.int32),
               "matches12": np.full((max(n // 2, 1), cap), -1, np.int32)}
        if full:
            out["kps"] = np.zeros((n, cap), KP_DTYPE)
            out["kps_un"] = np.zeros((n, cap), KP_DTYPE)
            out["desc"] = np.zeros((n, cap, 32), np.uint8)
        t = check(lib().orbx_stream_process_async(self.ex._h, ptr(imgs), n, self.nnratio, int(self.checkOri), self.window, chunk,
                                                  ptr(out.get("kps")), ptr(out.get("kps_un")), ptr(out.get("desc")), cap,
                                                  ptr(out["n"]), ptr(out["matches12"]), ptr(out["nmatches"]),
                                                  ptr(out["counters"])), self.ex._h)
        if not hasattr(self, "_flight"):
            self._flight = {}
        self._flight[t] = (imgs, out)   # the buffers stay alive until the step has been waited for
        return t

    def wait(self, ticket):
        check(lib().orbx_stream_wait(self.ex._h, ticket), self.ex._h)
        return self._flight.pop(ticket)[1]

    def features_in_area(self, slots, queries, cap=None):
        """Batched GetFeaturesInArea on the resident frames: queries [n, 5] = (x, y, r, minLevel, maxLevel), slots [n].
        -> list of index arrays in the reference's visiting order."""
        q = np.ascontiguousarray(queries, np.float32).reshape(-1, 5)
        sl = np.ascontiguousarray(slots, np.int32)
        cap = cap or self.ex.cap
        out = np.zeros((len(q), cap), np.int32)
        cnt = np.zeros(len(q), np.int32)
        check(lib().orbx_stream_features_in_area(self.ex._h, len(q), ptr(sl), ptr(q), ptr(out), cap, ptr(cnt)), self.ex._h)
        return [out[i, :min(int(cnt[i]), cap)].copy() for i in range(len(q))], cnt

    def score_models(self, n_pairs, H21=None, H12=None, F21=None, sigma=1.0):
        """CheckHomography / CheckFundamental for [n_pairs, n_hyp, 9] hypothesis sets on the resident pairs."""
        cap = self.ex.cap
        mats = {k: (None if M is None else np.ascontiguousarray(M, np.float32).reshape(n_pairs, -1, 9))
                for k, M in (("H21", H21), ("H12", H12), ("F21", F21))}
        nh = (mats["H21"] if mats["H21"] is not None else mats["F21"]).shape[1]
        out = {"n_list": np.zeros(n_pairs, np.int32)}
        for m in ("h", "f"):
            on = mats["H21" if m == "h" else "F21"] is not None
            out["scores_" + m] = np.zeros((n_pairs, nh), np.float32) if on else None
            out["best_" + m] = np.zeros(n_pairs, np.int32) if on else None
            out["inliers_" + m] = np.zeros((n_pairs, cap), np.uint8) if on else None
        check(lib().orbx_stream_score_models(self.ex._h, n_pairs, ptr(mats["H21"]), ptr(mats["H12"]), ptr(mats["F21"]), nh,
                                             sigma, ptr(out["scores_h"]), ptr(out["scores_f"]), ptr(out["best_h"]),
                                             ptr(out["best_f"]), ptr(out["inliers_h"]), ptr(out["inliers_f"]), cap,
                                             ptr(out["n_list"])), self.ex._h)
        return out
