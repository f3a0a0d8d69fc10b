"""The CPU oracle (oracle/orb_oracle.cpp) against the golden vectors in tests/golden/: outputs of the
reference's own sources (Features/ORBextractor.cpp, SlamTypes/Frame.cpp, Features/ORBmatcher.cpp,
Thirdparty/DBoW2/src/FORB.cpp compiled in place, tools/gen_golden.py) on the reference's bundled
images.  Bit-exact: keypoint records, descriptors, undistorted keys, bounds, grid, vnMatches12, count."""
import numpy as np
import pytest

from golden_util import CASES, LIGHT, Case, LightCase, same_records


def test_golden_present():
    assert len(CASES) >= 5


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_outputs(oracle, name):
    c = Case(name)
    E = oracle.Extractor(c.nf, 1.2, 8, c.ini, c.mn)
    z = c.z
    assert np.array_equal(E.scale, z["tab_scale"]) and np.array_equal(E.inv_scale, z["tab_inv_scale"])
    assert np.array_equal(E.sigma2, z["tab_sigma2"]) and np.array_equal(E.inv_sigma2, z["tab_inv_sigma2"])
    assert np.array_equal(E.per_level, z["tab_per_level"])
    bounds = oracle.image_bounds(c.imgs[0].shape[1], c.imgs[0].shape[0], c.K4, c.D4)
    assert np.array_equal(np.asarray(bounds), c.bounds)
    F = []
    for i, im in enumerate(c.imgs):
        k, d, mono = E(im)
        gk, gd, gku, goff, gidx = c.frame(i)
        assert same_records(k, gk), "keypoints of frame %d" % i
        assert np.array_equal(d, gd), "descriptors of frame %d" % i
        assert mono == len(gk)
        if i == 0:
            for l in (1, 7):
                assert int(E.level(l, True).astype(np.int64).sum()) == int(z["level%d_crc" % l])
        f = oracle.Frame(k, d, c.K4, c.D4, bounds)
        assert same_records(f.keys_un(), gku)
        off, idx = f.grid()
        assert np.array_equal(off, goff) and np.array_equal(idx, gidx)
        if c.lapping != (0, 0):
            k2, d2, mono2 = E(im, c.lapping)
            assert same_records(k2, z["lap_kps%d" % i]) and np.array_equal(d2, z["lap_desc%d" % i])
            assert mono2 == int(z["lap_mono%d" % i])
        F.append(f)
    for a in c.pairs():
        for w in c.windows:
            n, m12, _ = oracle.search_for_initialization(F[a], F[a + 1], 0.9, True, w)
            gn, gm = c.matches(a, w)
            assert n == gn and np.array_equal(m12, gm), "pair %d window %d" % (a, w)


def test_demo_pair_headline_counts():
    """config[0] as demo_initialization runs it: 2000 + 2000 keypoints, 129 matches (>= 100, so the demo proceeds)."""
    c = Case("demo_pair_th0")
    assert len(c.frame(0)[0]) == 2000 and len(c.frame(1)[0]) == 2000
    assert c.matches(0, 100)[0] == 129
    assert list(c.bounds) == [-78, 857, -61, 543]


@pytest.mark.parametrize("name", LIGHT)
def test_oracle_matches_reference_outputs_large_shapes(oracle, name):
    """KITTI-shape and 3840x2160 / 8000 keypoints, windows 100 and 8192 (all level-0 pairs)."""
    c = LightCase(name)
    E = oracle.Extractor(c.nf, 1.2, 8, 20, 7)
    bounds = oracle.image_bounds(c.W, c.H, c.K4, c.D4)
    F = []
    for i, im in enumerate(c.imgs):
        k, d, _ = E(im)
        f = oracle.Frame(k, d, c.K4, c.D4, bounds)
        c.check_frame(i, k, d, f.keys_un())
        F.append(f)
    for w in c.windows:
        n, m12, _ = oracle.search_for_initialization(F[0], F[1], 0.9, True, w)
        gn, gm = c.matches(w)
        assert n == gn and np.array_equal(m12, gm), w
