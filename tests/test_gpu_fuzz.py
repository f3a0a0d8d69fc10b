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


@pytest.mark.parametrize("cfg", _configs(8, seed=2026), ids=lambda c: "%dx%d_nf%d_sf%g_nl%d_th%d_%d" % c[:7])
def test_random_batches_both_policies(oracle, cfg):
    """The batched paths on random configurations (the committed slice of tools/gpu/fuzz_stream.py, which found the two defects of
    the ComputeKeyPointsOld path): a batch of 34..70 distinct-ish frames through orbx_stream_run and orbx_stream_process, then
    ComputeKeyPointsOld and a random lapping area through the single-frame path."""
    from orb_slam_tracking_b200 import frontend as fe, synth
    from orb_slam_tracking_b200.capi import OrbxError
    W, H, nf, sf, nl, ini, mn, dist, idx = cfg
    if W * H > 500 * 400:
        W, H = W // 2 + 80, H // 2 + 60
    nf = min(nf, 900)
    rng = np.random.default_rng(idx)
    P = int(rng.integers(17, 36))
    try:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H, max_batch=2 * P)
    except OrbxError:
        pytest.skip("shape outside the reference's range")
    R = oracle.Extractor(nf, sf, nl, ini, mn)
    K4 = np.array([0.8 * W, 0.8 * W, W / 2 - 3.5, H / 2 + 2.25], np.float32)
    D4 = np.array([-0.28, 0.07, 2e-4, 2e-5] if dist else [0, 0, 0, 0], np.float32)
    b = E.set_camera(K4, D4)
    base = synth.frames(8, W, H, first_pair=500 + 4 * idx)
    order = [(3 * i + (i // 8)) % 8 for i in range(2 * P)]
    imgs = np.ascontiguousarray(np.stack([base[j] for j in order]))
    ref = {}
    for j in set(order):
        k, d, _ = R(base[j])
        ref[j] = (k, d, oracle.Frame(k, d, K4, D4, b))
    win = int(rng.choice([100, 40, 4096]))
    S = fe.Stream(E, 0.9, True, win)
    for mode in ("run", "process"):
        if mode == "run":
            S.upload(imgs); S.run(2 * P); S.sync(); out = S.download(2 * P)
        else:
            out = S.process(imgs, chunk=int(rng.choice([0, 6, 10, 32])))
        for i, j in enumerate(order):
            rk, rd, _ = ref[j]
            n = int(out["n"][i])
            assert n == len(rk) and np.array_equal(out["kps"][i][:n].view(np.uint8), rk.view(np.uint8)) and np.array_equal(out["desc"][i][:n], rd), (mode, i)
        for p in range(0, P, 5):
            f1, f2 = ref[order[2 * p]][2], ref[order[2 * p + 1]][2]
            rn, rm12, rcnt = oracle.search_for_initialization(f1, f2, 0.9, True, win)
            assert int(out["nmatches"][p]) == rn and np.array_equal(out["matches12"][p][:f1.N], rm12) and np.array_equal(out["counters"][p], rcnt), (mode, p)
    lap = (int(rng.integers(0, W // 2)), int(rng.integers(W // 2, W)))
    for pol in (0, 1):
        try:
            E.set_keypoint_policy(pol)
        except OrbxError:
            continue   # the other policy refuses the shape (and so does the oracle: tests/test_old_policy.py)
        Rp = oracle.Extractor(nf, sf, nl, ini, mn, policy=pol)
        for im in base[:2]:
            rk, rd, rm = Rp(im, lapping=lap)
            k, d, m = E(im, vLappingArea=lap)
            assert m == rm and len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd), (pol, lap)
    E.close()
