"""CPU only: the oracle's OpenCV-primitive restatements (oracle/cv_prims.hpp) against the real OpenCV wheel on RANDOM sizes and
contents -- resize (any ratio between 1 and 3.2, both axes independent), reflect-101 frame, FAST with NMS on small and large images at
random thresholds, 7x7 Gaussian blur, fastAtan2, undistortPoints with random cameras, BGR / RGB to grey."""
import sys, numpy as np, cv2
sys.path.insert(0, '.')
from oracle import orb_oracle as orc
N, seed = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(seed)
bad = {}
def note(k): bad[k] = bad.get(k, 0) + 1
def img(h, w):
    kind = rng.integers(0, 3)
    a = rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == 1: a = cv2.GaussianBlur(a, (0, 0), float(rng.uniform(0.8, 3)))
    if kind == 2: a = (a > 128).astype(np.uint8) * np.uint8(rng.integers(30, 255))
    return np.ascontiguousarray(a)
for it in range(N):
    h, w = int(rng.integers(8, 420)), int(rng.integers(8, 560))
    a = img(h, w)
    dw, dh = max(1, int(round(w / rng.uniform(1.0, 3.2)))), max(1, int(round(h / rng.uniform(1.0, 3.2))))
    if not np.array_equal(cv2.resize(a, (dw, dh), interpolation=cv2.INTER_LINEAR), orc.resize_linear(a, dw, dh)): note("resize %dx%d->%dx%d" % (w, h, dw, dh))
    if h > 20 and w > 20 and not np.array_equal(cv2.copyMakeBorder(a, 19, 19, 19, 19, cv2.BORDER_REFLECT_101), orc.border101(a, 19)): note("border")
    th = int(rng.choice([0, 1, 7, 12, 20, 40, 100, 200, 254]))
    kp = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(a)
    xs, ys, sc = orc.fast(a, th)
    ref = [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kp]
    if ref != list(zip(xs.tolist(), ys.tolist(), sc.tolist())): note("fast th %d %dx%d" % (th, w, h))
    if h >= 4 and w >= 4 and not np.array_equal(cv2.GaussianBlur(a, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101), orc.blur7(a)): note("blur %dx%d" % (w, h))
    for _ in range(50):
        y, x = (float(v) for v in rng.integers(-200000, 200000, 2))
        if cv2.fastAtan2(y, x) != orc.fast_atan2(y, x): note("atan2")
    K4 = np.array([rng.uniform(200, 900), rng.uniform(200, 900), rng.uniform(100, 500), rng.uniform(100, 400)], np.float32)
    D4 = np.array([rng.uniform(-0.4, 0.2), rng.uniform(-0.1, 0.2), rng.uniform(-2e-3, 2e-3), rng.uniform(-2e-3, 2e-3)], np.float32)
    pts = rng.uniform(-50, 900, (60, 2)).astype(np.float32)
    K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
    if not np.array_equal(cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D4, None, None, K).reshape(-1, 2).view(np.uint32), orc.undistort(pts, K4, D4).view(np.uint32)): note("undistort")
    c = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if not np.array_equal(cv2.cvtColor(c, cv2.COLOR_BGR2GRAY), orc.bgr2gray(c)): note("bgr2gray")
print("iterations", N, "mismatches", bad or 0)
