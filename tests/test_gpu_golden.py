"""The CUDA path (through the C ABI) against the golden vectors in tests/golden/: outputs of the
reference's own sources on the reference's bundled images (tools/gen_golden.py).  Bit-exact:
level tables, keypoint records, descriptors, lapping split, undistorted keys, image bounds, the
64x48 grid, vnMatches12 and the returned count -- host-buffer calls and the device-resident stream."""
import numpy as np
import pytest

from golden_util import CASES, LIGHT, Case, LightCase, same_records

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_reference_outputs(name):
    from orb_slam_tracking_b200 import frontend as fe
    c = Case(name)
    H, W = c.imgs[0].shape
    E = fe.ORBextractor(c.nf, 1.2, 8, c.ini, c.mn, W, H, max_batch=c.n_frames)
    z = c.z
    assert np.array_equal(E.GetScaleFactors(), z["tab_scale"]) and np.array_equal(E.GetInverseScaleFactors(), z["tab_inv_scale"])
    assert np.array_equal(E.GetScaleSigmaSquares(), z["tab_sigma2"])
    assert np.array_equal(E.GetInverseScaleSigmaSquares(), z["tab_inv_sigma2"])
    assert np.array_equal(E.GetNumFeaturesPerLevel(), z["tab_per_level"])
    assert np.array_equal(E.set_camera(c.K4, c.D4), c.bounds)
    F = []
    for i, im in enumerate(c.imgs):
        gk, gd, gku, goff, gidx = c.frame(i)
        k, d, mono = E(im)
        assert same_records(k, gk), "keypoints of frame %d" % i
        assert np.array_equal(d, gd), "descriptors of frame %d" % i
        assert mono == len(gk)
        if i == 0:
            for l in (1, 7):
                assert int(E.pyramid_level(l, padded=True).astype(np.int64).sum()) == int(z["level%d_crc" % l])
        f = fe.Frame(E, kps=k, desc=d)
        assert same_records(f.mvKeysUn, gku)
        assert np.array_equal(f.grid_offsets, goff) and np.array_equal(f.grid_indices, gidx)
        if c.lapping != (0, 0):
            k2, d2, mono2 = E(im, vLappingArea=c.lapping)
            assert same_records(k2, z["lap_kps%d" % i]) and np.array_equal(d2, z["lap_desc%d" % i])
            assert mono2 == int(z["lap_mono%d" % i])
        F.append(f)
    M = fe.ORBmatcher(0.9, True)
    for a in c.pairs():
        for w in c.windows:
            n, m12, _ = M.SearchForInitialization(F[a], F[a + 1], w)
            gn, gm = c.matches(a, w)
            assert n == gn and np.array_equal(m12, gm), "pair %d window %d" % (a, w)
    # the device-resident stream: images in, match lists out
    for w in c.windows:
        S = fe.Stream(E, 0.9, True, w)
        nfr = c.n_frames - c.n_frames % 2
        S.upload(np.stack(c.imgs[:nfr]))
        S.run(nfr)
        out = S.download(nfr)
        for i in range(nfr):
            gk, gd, gku, _, _ = c.frame(i)
            n = int(out["n"][i])
            assert same_records(out["kps"][i, :n], gk) and np.array_equal(out["desc"][i, :n], gd)
            assert same_records(out["kps_un"][i, :n], gku)
        for a in c.pairs():
            gn, gm = c.matches(a, w)
            assert int(out["nmatches"][a // 2]) == gn and np.array_equal(out["matches12"][a // 2, :len(gm)], gm)
    E.close()


def test_color_ingest_matches_cvtcolor():
    """Row N1: Converter::toGray / cv::cvtColor on the device, fused into the ingest, against the
    real OpenCV wheel (BGR2GRAY and RGB2GRAY), odd width included; the extraction that follows equals the one
    on the host-converted gray image."""
    import cv2
    from orb_slam_tracking_b200 import frontend as fe
    rng = np.random.default_rng(11)
    for W, H in ((640, 480), (333, 250)):
        E = fe.ORBextractor(500, 1.2, 8, 20, 7, W, H, max_batch=2)
        E.set_camera(np.array([500, 500, W / 2, H / 2], np.float32), np.zeros(4, np.float32))
        bgr = rng.integers(0, 256, (2, H, W, 3), dtype=np.uint8)
        bgr[1] = cv2.GaussianBlur(bgr[1], (9, 9), 3)
        S = fe.Stream(E, 0.9, True, 100)
        for rgb, code in ((False, cv2.COLOR_BGR2GRAY), (True, cv2.COLOR_RGB2GRAY)):
            S.upload_color(bgr, rgb=rgb)
            S.run(2)
            out = S.download(2)
            for i in range(2):
                gray = cv2.cvtColor(bgr[i], code)
                assert np.array_equal(E.pyramid_level(0, slot=i), gray), (W, rgb, i)
                k, d, _ = E(gray)
                n = int(out["n"][i])
                assert n == len(k) and same_records(out["kps"][i, :n], k) and np.array_equal(out["desc"][i, :n], d)
        E.close()


@pytest.mark.parametrize("name", LIGHT)
def test_cuda_matches_reference_outputs_large_shapes(name):
    """BASELINE.json configs[3] / configs[4] against the reference's own outputs: 1241x376 / 2000 and 3840x2160 / 8000
    keypoints, SearchForInitialization with windows 100 and 8192 (brute force), every matcher path."""
    from orb_slam_tracking_b200 import frontend as fe
    c = LightCase(name)
    E = fe.ORBextractor(c.nf, 1.2, 8, 20, 7, c.W, c.H, max_batch=2)
    E.set_camera(c.K4, c.D4)
    F = []
    for i, im in enumerate(c.imgs):
        f = fe.Frame(E, im, fused=True)
        c.check_frame(i, f.mvKeys, f.mDescriptors, f.mvKeysUn)
        F.append(f)
    for w in c.windows:
        gn, gm = c.matches(w)
        for path in (0, 3, 1):
            E.set_option("match_path", path)
            n, m12, _ = fe.ORBmatcher(0.9, True).SearchForInitialization(F[0], F[1], w)
            assert n == gn and np.array_equal(m12, gm), (w, path)
    E.set_option("match_path", 0)
    S = fe.Stream(E, 0.9, True, 8192)
    S.upload(c.imgs)
    S.run(2)
    out = S.download(2)
    gn, gm = c.matches(8192)
    assert int(out["nmatches"][0]) == gn and np.array_equal(out["matches12"][0, :len(gm)], gm)
    E.close()
