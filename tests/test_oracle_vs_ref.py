"""The oracle restatement against the reference's own sources compiled in place
(oracle/_ref/libref_orb.so, built by oracle/build_ref.sh from /root/reference with the header
stand-in oracle/refshim).  Runs wherever the library exists (it is built in the container that
has /root/reference and travels with the snapshot)."""
import numpy as np
import pytest

from golden_util import same_records
from oracle import ref_orb as R
from orb_slam_tracking_b200 import synth

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref_orb.so not built (needs /root/reference)")

SHAPES = [(640, 480, 1000, 20, 7), (752, 480, 1200, 20, 7), (752, 480, 2000, 0, 0), (1241, 376, 2000, 20, 7),
          (320, 240, 500, 20, 7), (260, 250, 300, 12, 3)]


@pytest.mark.parametrize("W,H,nf,ini,mn", SHAPES)
def test_extract_frame_match(oracle, W, H, nf, ini, mn):
    K4, D4 = synth.camera_for(W, H)
    imgs = synth.frames(2, W, H, first_pair=11)
    Eo, Er = oracle.Extractor(nf, 1.2, 8, ini, mn), R.Extractor(nf, 1.2, 8, ini, mn)
    for a, b in zip(Er.tables(), (Eo.scale, Eo.inv_scale, Eo.sigma2, Eo.inv_sigma2, Eo.per_level)):
        assert np.array_equal(a, b)
    bounds = oracle.image_bounds(W, H, K4, D4)
    Fo, Fr = [], []
    for i in range(2):
        k, d, _ = Eo(imgs[i])
        fr = R.Frame(Er, imgs[i], K4, D4, first=(i == 0))
        rk, rku, rd = fr.get()
        assert same_records(k, rk) and np.array_equal(d, rd)
        for l in range(8):
            assert np.array_equal(Eo.level(l, True), Er.level(l, True))
        fo = oracle.Frame(k, d, K4, D4, bounds)
        assert np.array_equal(np.asarray(bounds), fr.bounds)
        assert same_records(fo.keys_un(), rku)
        go, gr = fo.grid(), fr.grid()
        assert np.array_equal(go[0], gr[0]) and np.array_equal(go[1], gr[1])
        rng = np.random.default_rng(5)
        for _ in range(20):
            x, y, r = float(rng.uniform(-30, W + 30)), float(rng.uniform(-30, H + 30)), float(rng.uniform(5, 150))
            lv = (-1, -1) if rng.random() < 0.5 else (0, int(rng.integers(0, 4)))
            assert np.array_equal(fo.features_in_area(x, y, r, *lv), fr.features_in_area(x, y, r, *lv))
        Fo.append(fo)
        Fr.append(fr)
    for w in (100, 40, 2000):
        for ratio, ori in ((0.9, True), (0.6, False)):
            n, m12, _ = oracle.search_for_initialization(Fo[0], Fo[1], ratio, ori, w)
            rn, rm = R.search_for_initialization(Fr[0], Fr[1], ratio, ori, w)
            assert n == rn and np.array_equal(m12, rm)


def test_lapping_area_split(oracle):
    W, H, nf = 640, 480, 800
    im = synth.frames(2, W, H, first_pair=2)[0]
    Eo, Er = oracle.Extractor(nf, 1.2, 8, 20, 7), R.Extractor(nf, 1.2, 8, 20, 7)
    for lap in ((0, 0), (100, 300), (0, 639), (630, 640)):
        k, d, m = Eo(im, lap)
        rk, rd, rm = Er(im, lap)
        assert m == rm and same_records(k, rk) and np.array_equal(d, rd)


def test_degenerate_images(oracle):
    W, H = 640, 480
    Eo, Er = oracle.Extractor(500, 1.2, 8, 20, 7), R.Extractor(500, 1.2, 8, 20, 7)
    flat = np.full((H, W), 90, np.uint8)
    one = flat.copy()
    one[200:260, 300:380] = 200
    noise = np.random.default_rng(1).integers(0, 256, (H, W), dtype=np.uint8)
    for im in (flat, one, noise):
        k, d, m = Eo(im)
        rk, rd, rm = Er(im)
        assert m == rm and same_records(k, rk) and np.array_equal(d, rd)


def test_forb_distance(oracle):
    rng = np.random.default_rng(0)
    for _ in range(200):
        a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
        assert R.forb_distance(a, b) == float(oracle.descriptor_distance(a, b))
