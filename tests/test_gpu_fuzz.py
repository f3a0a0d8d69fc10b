"""Parity on RANDOM configurations: image shape (odd widths, widths that are no multiple of 4 or 16, tall and wide aspect ratios inside
the range the reference handles), number of features, scale factor, number of levels, thresholds, distortion.  For each configuration
one frame pair goes through extraction, the Frame glue and SearchForInitialization on the GPU (C ABI) and through the oracle; keypoint
records, descriptors, undistorted keys, grid and vnMatches12 must be bit-equal.  The seed is fixed: the configurations are the same
on every run."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _configs(n, seed=20261105):
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n:
        W, H = int(rng.integers(160, 1000)), int(rng.integers(120, 720))
        if not 0.5 <= W / H < 8.4:      # nIni = round(W' / H') must be >= 1 (the reference divides by it) and the roots fit the key
            continue
        sf = float(rng.choice([1.1, 1.2, 1.2, 1.25, 1.3, 1.5, 2.0]))
        nl = int(rng.integers(2, 9))
        # the smallest level must still hold the 16-pixel border on both sides and one FAST cell
        if min(W, H) / sf ** (nl - 1) < 70:
            continue
        nf = int(rng.integers(60, 2500))
        ini = int(rng.choice([20, 20, 7, 12, 31, 40, 0]))
        mn = int(rng.choice([7, 7, 3, 0, 12, ini]))
        dist = rng.random() < 0.5
        out.append((W, H, nf, sf, nl, ini, mn, dist, len(out)))
    return out


@pytest.mark.parametrize("cfg", _configs(32), ids=lambda c: "%dx%d_nf%d_sf%g_nl%d_th%d_%d" % c[:7])
def test_random_configuration(oracle, cfg):
    from orb_slam_tracking_b200 import frontend as fe, synth
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    from orb_slam_tracking_b200.capi import OrbxError, ERR_ASPECT
    R = oracle.Extractor(nf, sf, nl, ini, mn)
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H)
    except OrbxError as err:
        # a level whose bordered aspect ratio rounds to 0 quadtree roots: the reference indexes an empty vector
        # (Features/ORBextractor.cpp:706).  orbx_create refuses the shape; the oracle must refuse the image for the same reason
        assert err.code == ERR_ASPECT
        with pytest.raises(RuntimeError, match="status -2"):
            R(synth.frame_pair(100 + idx, W, H)[0])
        return
    bounds = E.set_camera(K4, D4)
    assert np.array_equal(bounds, oracle.image_bounds(W, H, K4, D4))
    G, O = [], []
    for im in synth.frame_pair(100 + idx, W, H):
        rk, rd, rmono = R(im)
        k, d, mono = E(im)
        assert len(k) == len(rk) and mono == rmono, (len(k), len(rk))
        assert np.array_equal(k.view(np.uint8), rk.view(np.uint8)), "keypoint records"
        assert np.array_equal(d, rd), "descriptors"
        F = fe.Frame(E, im)
        Fo = oracle.Frame(rk, rd, K4, D4, bounds)
        assert F.N == Fo.N
        assert np.array_equal(F.mvKeysUn.view(np.uint8), Fo.keys_un().view(np.uint8)), "undistorted keys"
        off, gidx = Fo.grid()
        assert np.array_equal(F.grid_offsets, off) and np.array_equal(F.grid_indices, gidx), "grid"
        G.append(F); O.append(Fo)
    for nnratio, ori, window in ((0.9, True, 100), (0.7, False, 40), (0.9, True, 4096)):
        n, m12, cnt = fe.ORBmatcher(nnratio, ori).SearchForInitialization(G[0], G[1], window)
        rn, rm12, rcnt = oracle.search_for_initialization(O[0], O[1], nnratio, ori, window)
        assert n == rn and np.array_equal(m12, rm12) and np.array_equal(cnt, rcnt), (nnratio, ori, window)
    E.close()
