"""Pins the oracle's OpenCV-primitive restatements (oracle/cv_prims.hpp) bit-for-bit against
the real OpenCV 4.13.0 wheel.  The reference calls these primitives at
Features/ORBextractor.cpp:158,1109,1119,1601,1676,1689,1708 and SlamTypes/Frame.cpp:119,150.
CPU-only; skipped if cv2 is not importable."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


def _rand_img(rng, h, w, smooth=True):
    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    if smooth:
        img = cv2.GaussianBlur(img, (0, 0), 1.5)
        img = cv2.normalize(img, None, 0, 255, cv2.NORM_MINMAX)
    return np.ascontiguousarray(img)


@pytest.mark.parametrize("src,dst", [((480, 640), (400, 533)), ((400, 533), (333, 444)), ((161, 214), (134, 179)),
                                     ((376, 1241), (313, 1034)), ((97, 131), (80, 111)), ((64, 64), (77, 90)),
                                     ((480, 752), (400, 627)), ((51, 67), (17, 23))])
def test_resize_linear(oracle, src, dst):
    rng = np.random.default_rng(1)
    for smooth in (True, False):
        img = _rand_img(rng, *src, smooth=smooth)
        ref = cv2.resize(img, (dst[1], dst[0]), interpolation=cv2.INTER_LINEAR)
        got = oracle.resize_linear(img, dst[1], dst[0])
        assert np.array_equal(ref, got)


def test_resize_pyramid_chain(oracle):
    rng = np.random.default_rng(2)
    img = _rand_img(rng, 480, 640)
    E = oracle.Extractor(1000, 1.2, 8, 20, 7)
    E(img)
    prev = img
    for l in range(1, 8):
        w, h = E.level_size(l)
        ref = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(ref, E.level(l)), l
        pad = cv2.copyMakeBorder(ref, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        assert np.array_equal(pad, E.level(l, padded=True)), l
        prev = ref


def test_border101(oracle):
    rng = np.random.default_rng(3)
    for shape in ((30, 41), (134, 179), (20, 25)):
        img = _rand_img(rng, *shape, smooth=False)
        ref = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        assert np.array_equal(ref, oracle.border101(img, 19))


@pytest.mark.parametrize("th", [0, 1, 7, 20, 60])
def test_fast_nms(oracle, th):
    rng = np.random.default_rng(4 + th)
    det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    for shape in ((41, 41), (36, 47), (7, 7), (8, 30), (6, 40), (120, 160), (47, 36)):
        for smooth in (True, False):
            img = _rand_img(rng, *shape, smooth=smooth)
            kps = det.detect(img)
            xs, ys, sc = oracle.fast(img, th)
            assert len(kps) == len(xs), (shape, smooth)
            for k, x, y, s in zip(kps, xs, ys, sc):
                assert (k.pt[0], k.pt[1], k.response) == (x, y, s)
                assert k.size == 7.0 and k.angle == -1.0 and k.octave == 0 and k.class_id == -1


def test_fast_on_noncontiguous_roi(oracle):
    rng = np.random.default_rng(9)
    img = _rand_img(rng, 200, 300)
    roi = img[16:16 + 41, 51:51 + 41]
    det = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(np.ascontiguousarray(roi))
    import ctypes as C
    xs, ys, sc = (np.empty(4096, np.int32) for _ in range(3))
    base = img.ctypes.data + 16 * img.strides[0] + 51
    n = oracle.lib().orc_fast(C.c_void_p(base), 41, 41, img.strides[0], 7, xs.ctypes.data_as(C.c_void_p),
                              ys.ctypes.data_as(C.c_void_p), sc.ctypes.data_as(C.c_void_p), 4096)
    assert n == len(kps)
    assert [(k.pt[0], k.pt[1], k.response) for k in kps] == [(x, y, s) for x, y, s in zip(xs[:n], ys[:n], sc[:n])]


def test_fast_atan2(oracle):
    rng = np.random.default_rng(5)
    ys = rng.integers(-60000, 60001, 20000)
    xs = rng.integers(-60000, 60001, 20000)
    for y, x in list(zip(ys, xs)) + [(0, 0), (0, 5), (5, 0), (0, -5), (-5, 0), (7, 7), (-7, 7), (1, 60000)]:
        assert cv2.fastAtan2(float(y), float(x)) == oracle.fast_atan2(float(y), float(x)), (y, x)


def test_gaussian_blur(oracle):
    rng = np.random.default_rng(6)
    for shape in ((480, 640), (134, 179), (97, 131), (20, 9), (7, 7)):
        for smooth in (True, False):
            img = _rand_img(rng, *shape, smooth=smooth)
            ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            assert np.array_equal(ref, oracle.blur7(img)), shape


def test_undistort_points(oracle):
    rng = np.random.default_rng(7)
    K4 = np.array([609.2855, 609.3422, 351.4274, 237.7324], np.float32)
    for D4 in (np.array([-0.3492, 0.1363, 0, 0], np.float32), np.array([-0.28, 0.07, 1e-3, -2e-4], np.float32)):
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        pts = np.stack([rng.uniform(-5, 757, 5000), rng.uniform(-5, 485, 5000)], 1).astype(np.float32)
        pts[:4] = [[0, 0], [752, 0], [0, 480], [752, 480]]
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D4, None, None, K).reshape(-1, 2)
        got = oracle.undistort(pts, K4, D4)
        assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))
    b = oracle.image_bounds(752, 480, K4, np.array([-0.3492, 0.1363, 0, 0], np.float32))
    assert list(b) == [-78, 857, -61, 543]


def test_bgr2gray(oracle):
    rng = np.random.default_rng(8)
    bgr = rng.integers(0, 256, (77, 91, 3), dtype=np.uint8)
    assert np.array_equal(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), oracle.bgr2gray(bgr))


def test_descriptor_distance(oracle):
    rng = np.random.default_rng(10)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8)
        b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert oracle.descriptor_distance(a, b) == int(np.unpackbits(a ^ b).sum())
        assert oracle.descriptor_distance(a, b) == int(cv2.norm(a, b, cv2.NORM_HAMMING))


def test_retain_best_selection_and_order():
    """cv::KeyPointsFilter::retainBest (ComputeKeyPointsOld, Features/ORBextractor.cpp:1452,1482) is not exposed by the wheel, but
    cv::ORB runs it: with one level and FAST_SCORE, ORB::detect returns retainBest(runByImageBorder(FAST keypoints), nfeatures)
    untouched (modules/features2d/src/orb.cpp, computeKeyPoints).  The restated retain_best (oracle/cv_prims.hpp: nth_element by
    response, the ties of the n-th kept, libstdc++ element order) must reproduce that list: same keypoints, same ORDER."""
    from oracle import orb_oracle as orc
    rng = np.random.default_rng(0)
    for trial in range(6):
        h, w = [(240, 320), (480, 640), (200, 500)][trial % 3]
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        img = cv2.normalize(cv2.GaussianBlur(img, (0, 0), 1.0 + 0.3 * trial), None, 0, 255, cv2.NORM_MINMAX)
        t = [20, 7, 12][trial % 3]
        b = 31
        kb = [kp for kp in cv2.FastFeatureDetector_create(t, True).detect(img) if b <= kp.pt[0] < w - b and b <= kp.pt[1] < h - b]
        resp = np.array([kp.response for kp in kb], np.float32)
        assert len(kb) > 500
        for n in (1, 7, 50, 200, 1000, len(kb) - 1, len(kb), len(kb) + 5):
            orb = cv2.ORB_create(nfeatures=n, scaleFactor=1.2, nlevels=1, edgeThreshold=b, firstLevel=0, WTA_K=2,
                                 scoreType=cv2.ORB_FAST_SCORE, patchSize=31, fastThreshold=t)
            ref = [(kp.pt[0], kp.pt[1], kp.response) for kp in orb.detect(img)]
            got = [(kb[i].pt[0], kb[i].pt[1], kb[i].response) for i in orc.retain_best(resp, n)]
            assert got == ref, (trial, n)


def test_undistort_points_negative_radial_factor(oracle):
    """OpenCV 4.x ends the fixed-point iteration at the start value when the radial factor turns negative (points far outside the
    distortion model's range; its regression test 14583).  Found by tools/fuzz_oracle_vs_cv2.py with random cameras."""
    rng = np.random.default_rng(5)
    for K4, D4 in (((250, 250, 320, 240), (-0.45, 0.0, 1e-3, -1e-3)), ((200, 300, 100, 400), (-0.4, 0.2, -2e-3, 2e-3)), ((609.2855, 609.3422, 351.4274, 237.7324), (-0.3492, 0.1363, 0, 0))):
        K4, D4 = np.array(K4, np.float32), np.array(D4, np.float32)
        pts = rng.uniform(-50, 900, (400, 2)).astype(np.float32)
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D4, None, None, K).reshape(-1, 2)
        assert np.array_equal(ref.view(np.uint32), oracle.undistort(pts, K4, D4).view(np.uint32))
